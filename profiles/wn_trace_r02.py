"""Per-phase timing of the round-2 fused WN-layer kernel (per-CTA globaltimer stamps, -DGTK_INSTRUMENT build) on the bench
workload.  Run with GTK_LIB pointing at the instrumented library:
    GTK_EXTRA_CFLAGS=-DGTK_INSTRUMENT python -m glowtts_b200.build --force   (then copy lib/ aside and rebuild the product)
    GTK_LIB=profiles/_instr/libglowtts_b200.so python profiles/wn_trace_r02.py
Stamps (thread 0 of the first epilogue warp): start, setup done, GEMM-1 chunk 0 complete, last gate epilogue done,
GEMM 2 complete, epilogue 2 done."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from glowtts_b200 import _lib
from glowtts_b200.hot_path import sequence_mask

run = bench.Runner(torch.device("cuda", 0), 1, "bf16")
for _ in range(3):
    run.step_resident()
torch.cuda.synchronize()
y, y_len, x_m, x_len = run.dev
mask = sequence_mask(y_len, y.shape[2]).float().unsqueeze(1)
names = ["setup", "chunk0 MMA", "rest of GEMM1 + gate epilogues", "GEMM2", "epilogue 2"]
n_ctas = 128
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
    t = t[t[:, 0] > t[:, 0].max() - 200_000]        # CTAs of the last launch only (dead tiles keep stale stamps)
    t = t[(np.diff(t, axis=1) >= 0).all(axis=1)]
    t0 = t[:, 0].min()
    d = np.diff(t, axis=1) / 1e3
    print(f"--- {mode}: {len(t)} live CTAs, last layer of the last block")
    print(f"kernel span (first CTA start -> last CTA end) {(t[:, 5].max() - t0) / 1e3:7.2f} us;"
          f" CTA start spread {(t[:, 0].max() - t0) / 1e3:.2f} us")
    print(f"{'phase':32s} {'min':>7s} {'median':>7s} {'max':>7s}  (us per CTA)")
    for i, n in enumerate(names):
        print(f"{n:32s} {d[:, i].min():7.2f} {np.median(d[:, i]):7.2f} {d[:, i].max():7.2f}")
    tot = (t[:, 5] - t[:, 0]) / 1e3
    print(f"{'total':32s} {tot.min():7.2f} {np.median(tot):7.2f} {tot.max():7.2f}")
    wb = (ctypes.c_longlong * (n_ctas * 4))()
    _lib.call("gtk_debug_wn_wait", wb, n_ctas)
    w = np.array(list(wb), dtype=np.int64).reshape(n_ctas, 4)[:len(t)]
    print("SM cycles (median over CTAs): MMA thread waiting for weight stages %d of %d in GEMM 1; producer waiting for free "
          "stages %d of %d" % tuple(np.median(w, axis=0)))
