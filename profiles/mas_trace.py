"""DP-warp phase breakdown (globaltimer stamps) of the TMA-path MAS kernel.  usage: python profiles/mas_trace.py [1|5]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from glowtts_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda", 0)
for which in sys.argv[1:] or ["1", "5"]:
    B, TX, TY = (32, 180, 860) if which == "1" else (512, 400, 2000)
    gen = torch.Generator(device=dev).manual_seed(1234)
    value = torch.randn(B, TX, TY, device=dev, generator=gen)
    t_xs = torch.full((B,), TX, device=dev, dtype=torch.int32)
    t_ys = torch.full((B,), TY, device=dev, dtype=torch.int32)
    for _ in range(3):
        ops.mas_path(value, t_xs, t_ys, 0)
    torch.cuda.synchronize()
    n = min(B, 1024)
    buf = (ctypes.c_ulonglong * (n * 6))()
    _lib.call("gtk_debug_mas_trace", buf, n)
    t = np.array(list(buf), dtype=np.int64).reshape(n, 6)
    d = np.diff(t, axis=1) / 1e3
    print(f"MAS B={B} T_x={TX} T_y={TY}: kernel span {(t[:, 5].max() - t[:, 0].min()) / 1e3:.2f} us;"
          f" CTA start spread {(t[:, 0].max() - t[:, 0].min()) / 1e3:.2f} us")
    print(f"{'phase':16s} {'min':>8s} {'median':>8s} {'max':>8s}  (us per CTA)")
    for i, name in enumerate(["first tile", "DP", "backtrack", "barrier", "emit"]):
        print(f"{name:16s} {d[:, i].min():8.2f} {np.median(d[:, i]):8.2f} {d[:, i].max():8.2f}")
    wbuf = (ctypes.c_longlong * n)()
    _lib.call("gtk_debug_mas_wait", wbuf, n)
    w = np.array(list(wbuf), dtype=np.int64)
    print(f"DP warp waiting for tiles (after the first): median {np.median(w)} SM cycles = {np.median(w) / 1.965e3:.2f} us")
    print(f"{'CTA total':16s} {(t[:, 5] - t[:, 0]).min() / 1e3:8.2f} {np.median(t[:, 5] - t[:, 0]) / 1e3:8.2f} {(t[:, 5] - t[:, 0]).max() / 1e3:8.2f}")
