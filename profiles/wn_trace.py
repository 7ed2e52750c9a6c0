"""Per-phase timing of the fused WN-layer kernel (CTA (0,0) globaltimer stamps) on the bench workload."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from glowtts_b200 import _lib
run = bench.Runner(torch.device("cuda", 0), 1, "bf16")
for _ in range(3):
    run.step_resident()
torch.cuda.synchronize()
from glowtts_b200.hot_path import sequence_mask
y, y_len, x_m, x_len = run.dev
mask = sequence_mask(y_len, y.shape[2]).float().unsqueeze(1)
names = ["setup", "GEMM1", "epi1", "GEMM2", "epi2"]
import numpy as np
T2 = y.shape[2] // 2
n_ctas = ((T2 + 127) // 128) * y.shape[0]
for mode in ("inference", "training"):
    if mode == "inference":
        with torch.no_grad():
            run.dec(y, mask)
    else:
        run.dec(y, mask)          # parameters require grad -> activations are saved
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * (n_ctas * 6))()
    _lib.call("gtk_debug_wn_trace", buf, n_ctas)
    t = np.array(list(buf), dtype=np.int64).reshape(n_ctas, 6)
    t0 = t[:, 0].min()
    d = np.diff(t, axis=1) / 1e3
    print(f"--- {mode}: {n_ctas} CTAs, last layer of the last block")
    print(f"kernel span (first CTA start -> last CTA end) {(t[:, 5].max() - t0) / 1e3:7.2f} us;"
          f" CTA start spread {(t[:, 0].max() - t0) / 1e3:.2f} us")
    print(f"{'phase':8s} {'min':>7s} {'median':>7s} {'max':>7s}  (us per CTA)")
    for i, n in enumerate(names):
        print(f"{n:8s} {d[:, i].min():7.2f} {np.median(d[:, i]):7.2f} {d[:, i].max():7.2f}")
    tot = (t[:, 5] - t[:, 0]) / 1e3
    print(f"{'total':8s} {tot.min():7.2f} {np.median(tot):7.2f} {tot.max():7.2f}")
