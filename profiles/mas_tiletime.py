"""Per-tile cycle split of the MAS DP warps (build mas.cu with -DGTK_MAS_EXP_TILETIME, load it through GTK_LIB):
tile 10 of utterance 0, DP warps 0..3: [wait value tile, mailbox in, 32 columns, bits store + arrives]."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from glowtts_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda", 0)
B, TX, TY = 32, 180, 860
gen = torch.Generator(device=dev).manual_seed(1234)
value = torch.randn(B, TX, TY, device=dev, generator=gen)
t_xs = torch.full((B,), TX, device=dev, dtype=torch.int32)
t_ys = torch.full((B,), TY, device=dev, dtype=torch.int32)
for _ in range(3):
    ops.mas_path(value, t_xs, t_ys, 0)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * (1024 * 6))()
_lib.call("gtk_debug_mas_trace", buf, 1024)
t = np.array(list(buf), dtype=np.int64).reshape(1024, 6)
for w in range(4):
    r = t[900 + w]
    print(f"DP warp {w}: wait tile {r[0]:6d}  mailbox {r[1]:6d}  columns {r[2]:6d}  store+arrive {r[3]:6d}  (SM cycles), tile start clock {r[4]}")
