"""One MAS call per BASELINE config (1: B=32 180x860, 5: B=512 400x2000) for ncu."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from glowtts_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
which = sys.argv[1] if len(sys.argv) > 1 else "1"
B, TX, TY = (32, 180, 860) if which == "1" else (int(sys.argv[2]) if len(sys.argv) > 2 else 512, 400, 2000)
gen = torch.Generator(device=dev).manual_seed(1234)
value = torch.randn(B, TX, TY, device=dev, generator=gen)
t_xs = torch.full((B,), TX, device=dev, dtype=torch.int32)
t_ys = torch.full((B,), TY, device=dev, dtype=torch.int32)
for _ in range(3):
    out = ops.mas_path(value, t_xs, t_ys, 0)
torch.cuda.synchronize()
print("ok", float(out[0].sum()))
