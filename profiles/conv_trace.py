"""Per-CTA phase breakdown of the tcgen05 conv kernel (globaltimer stamps) for each epilogue/tap variant of the step."""
import ctypes, os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
import torch
import bench
from glowtts_b200 import _lib

run = bench.Runner(torch.device("cuda", 0), 1, "bf16")
for _ in range(2):
    run.step_resident()
torch.cuda.synchronize()
n_ctas = 4 * 32
names = ["setup", "mainloop", "epilogue", "stores"]
for label, epi, taps in (("in_layer data gradient (plain, k=5; last launch = layer 0 of block 0)", 0, 5),
                         ("gate backward + res_skip data gradient (k=1)", 1, 1),
                         ("end conv + affine coupling (k=1)", 2, 1),
                         ("plain k=1 (last launch = start-conv data gradient of block 0)", 0, 1)):
    _lib.call("gtk_debug_conv_trace_select", epi, taps)
    run.step_resident()
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * (n_ctas * 5))()
    _lib.call("gtk_debug_conv_trace", buf, n_ctas)
    t = np.array(list(buf), dtype=np.int64).reshape(n_ctas, 5)
    d = np.diff(t, axis=1) / 1e3
    print(f"--- {label}")
    print(f"kernel span {(t[:, 4].max() - t[:, 0].min()) / 1e3:7.2f} us; CTA start spread {(t[:, 0].max() - t[:, 0].min()) / 1e3:.2f} us")
    print(f"{'phase':9s} {'min':>7s} {'median':>7s} {'max':>7s}  (us per CTA)")
    for i, n in enumerate(names):
        print(f"{n:9s} {d[:, i].min():7.2f} {np.median(d[:, i]):7.2f} {d[:, i].max():7.2f}")
    wbuf = (ctypes.c_longlong * (n_ctas * 2))()
    _lib.call("gtk_debug_conv_wait", wbuf, n_ctas)
    w = np.array(list(wbuf), dtype=np.int64).reshape(n_ctas, 2) / 1.965e3      # cycles -> us at 1965 MHz
    print(f"mainloop waits (median): producer on a free stage {np.median(w[:, 0]):.2f} us, MMA thread on data {np.median(w[:, 1]):.2f} us")
_lib.call("gtk_debug_conv_trace_select", -1, -1)
