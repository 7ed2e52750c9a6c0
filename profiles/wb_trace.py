"""Per-CTA phase breakdown (globaltimer stamps) of the fused WaveNet-layer backward kernel: last launch of a step."""
import ctypes, os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
import torch
import bench
from glowtts_b200 import _lib

run = bench.Runner(torch.device("cuda", 0), 1, "bf16")
for _ in range(3):
    run.step_resident()
torch.cuda.synchronize()
n_ctas = 4 * 32
buf = (ctypes.c_ulonglong * (n_ctas * 7))()
_lib.call("gtk_debug_wb_trace", buf, n_ctas)
t = np.array(list(buf), dtype=np.int64).reshape(n_ctas, 7)
d = np.diff(t, axis=1) / 1e3
print(f"fused layer backward (layer 0 of block 0): kernel span {(t[:, 6].max() - t[:, 0].min()) / 1e3:.2f} us;"
      f" CTA start spread {(t[:, 0].max() - t[:, 0].min()) / 1e3:.2f} us")
print(f"{'phase':12s} {'min':>7s} {'median':>7s} {'max':>7s}  (us per CTA)")
for i, n in enumerate(["setup", "GEMM a", "epilogue a", "GEMM b", "epilogue b", "stores"]):
    print(f"{n:12s} {d[:, i].min():7.2f} {np.median(d[:, i]):7.2f} {d[:, i].max():7.2f}")
