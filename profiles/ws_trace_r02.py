"""Per-phase timing of the whole-WaveNet kernel (csrc/wn_stack.cu) on the bench workload: per-CTA globaltimer stamps of
the LAST launch (last block), -DGTK_INSTRUMENT build.
    GTK_EXTRA_CFLAGS=-DGTK_INSTRUMENT GTK_LIBDIR=profiles/_instr GTK_OBJDIR=profiles/_instr/build python -m glowtts_b200.build
    GTK_LIB=profiles/_instr/libglowtts_b200.so python profiles/ws_trace_r02.py"""
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
L, N, n_ctas = 4, 24, 160
for mode in ("inference", "training"):
    if mode == "inference":
        with torch.no_grad():
            run.dec(y, mask)
    else:
        run.dec(y, mask)
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * (n_ctas * N))()
    _lib.call("gtk_debug_ws_trace", buf, n_ctas)
    t = np.array(list(buf), dtype=np.int64).reshape(n_ctas, N)[:, :4 + 4 * L]
    t = t[t[:, 0] > t[:, 0].max() - 400_000]
    t = t[(np.diff(t, axis=1) >= 0).all(axis=1)]
    t0 = t[:, 0].min()
    d = np.diff(t, axis=1) / 1e3
    names = ["prologue (incl. griddepcontrol.wait)", "start conv + epilogue"]
    for l in range(L):
        names += [f"L{l} GEMM 1 (incl. tile_ready wait)", f"L{l} epilogue 1", f"L{l} GEMM 2", f"L{l} epilogue 2"]
    names += ["final store"]
    print(f"--- {mode}: {len(t)} live CTAs; kernel span {(t[:, -1].max() - t0) / 1e3:.2f} us, CTA start spread {(t[:, 0].max() - t0) / 1e3:.2f} us")
    print(f"{'phase':40s} {'min':>7s} {'median':>7s} {'max':>7s}  (us per CTA)")
    for i, n in enumerate(names):
        print(f"{n:40s} {d[:, i].min():7.2f} {np.median(d[:, i]):7.2f} {d[:, i].max():7.2f}")
    tot = (t[:, -1] - t[:, 0]) / 1e3
    print(f"{'total':40s} {tot.min():7.2f} {np.median(tot):7.2f} {tot.max():7.2f}")
    wb = (ctypes.c_longlong * (n_ctas * 8))()
    _lib.call("gtk_debug_ws_wait", wb, n_ctas)
    w = np.array(list(wb), dtype=np.int64).reshape(n_ctas, 8)[:len(t)]
    print("GEMM 1 per layer, SM cycles (median): waiting for weight stages", np.median(w[:, :4], axis=0), "of", np.median(w[:, 4:], axis=0))
