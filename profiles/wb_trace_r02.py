"""Per-CTA phase breakdown (globaltimer stamps) + ring wait counters of the fused WaveNet-layer backward kernel and of the
forward kernel: last launches of a training step.  Needs the -DGTK_INSTRUMENT library (GTK_LIB=...)."""
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
n_ctas = 128


def live(t):
    t = t[t[:, 0] > t[:, 0].max() - 200_000]
    return t[(np.diff(t, axis=1) >= 0).all(axis=1)]


buf = (ctypes.c_ulonglong * (n_ctas * 7))()
_lib.call("gtk_debug_wb_trace", buf, n_ctas)
t = live(np.array(list(buf), dtype=np.int64).reshape(n_ctas, 7))
d = np.diff(t, axis=1) / 1e3
print(f"fused layer backward (layer 0 of block 0): {len(t)} live CTAs, kernel span {(t[:, 6].max() - t[:, 0].min()) / 1e3:.2f} us")
print(f"{'phase':12s} {'min':>7s} {'median':>7s} {'max':>7s}  (us per CTA)")
for i, n in enumerate(["setup", "GEMM a", "epilogue a", "GEMM b", "epilogue b", "stores"]):
    print(f"{n:12s} {d[:, i].min():7.2f} {np.median(d[:, i]):7.2f} {d[:, i].max():7.2f}")
wb = (ctypes.c_longlong * (n_ctas * 4))()
_lib.call("gtk_debug_wb_wait", wb, n_ctas)
w = np.array(list(wb), dtype=np.int64).reshape(n_ctas, 4)[:len(t)]
print("GEMM b, SM cycles (median): MMA thread waiting for weight stages %d of %d; producer waiting for free stages %d of %d"
      % tuple(np.median(w, axis=0)))

buf = (ctypes.c_ulonglong * (n_ctas * 6))()
_lib.call("gtk_debug_wn_trace", buf, n_ctas)
t = live(np.array(list(buf), dtype=np.int64).reshape(n_ctas, 6))
d = np.diff(t, axis=1) / 1e3
print(f"fused layer forward (training, last layer of the last block): {len(t)} live CTAs")
for i, n in enumerate(["setup", "GEMM 1", "epilogue 1", "GEMM 2", "epilogue 2"]):
    print(f"{n:12s} {d[:, i].min():7.2f} {np.median(d[:, i]):7.2f} {d[:, i].max():7.2f}")
wb = (ctypes.c_longlong * (n_ctas * 4))()
_lib.call("gtk_debug_wn_wait", wb, n_ctas)
w = np.array(list(wb), dtype=np.int64).reshape(n_ctas, 4)[:len(t)]
print("GEMM 1, SM cycles (median): MMA thread waiting for operand stages %d of %d; producer waiting for free stages %d of %d"
      % tuple(np.median(w, axis=0)))
