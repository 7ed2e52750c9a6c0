"""Does a row-shifted (non atom-aligned) K-major SWIZZLE_128B UMMA descriptor address the rows TMA wrote?"""
import os, sys
sys.path.insert(0, os.getcwd())
import torch
from glowtts_b200 import _lib
torch.manual_seed(0)
a = torch.randn(136, 64, device="cuda").to(torch.bfloat16)
b = torch.randn(64, 64, device="cuda").to(torch.bfloat16)
out = torch.zeros(8, 128, 64, device="cuda")
_lib.call("gtk_debug_shift_probe", a.data_ptr(), b.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
for s in range(1, 5):
    ref = a[s:s + 128].float() @ b.float().t()
    for mode in range(2):
        got = out[(s - 1) * 2 + mode]
        err = (got - ref).abs().max().item()
        # which shift does the result actually correspond to?
        best = min(range(0, 9), key=lambda q: (got - a[q:q + 128].float() @ b.float().t()).abs().max().item())
        print(f"shift {s} base_offset mode {mode}: max |err| {err:.3e}  (closest to shift {best})")
