"""Dead time between two dependent kernels of the forward chain: last CTA of the last fused WN layer kernel (block 11)
finishing -> CTAs of the following `end` conv (affine epilogue) leaving griddepcontrol.wait.  Instrumented library."""
import ctypes, os, sys
sys.path.insert(0, os.getcwd())
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
_lib.call("gtk_debug_conv_trace_select", 2, 1)
for pdl in ("graph-less eager",):
    with torch.no_grad():
        run.dec(y, mask)
    torch.cuda.synchronize()
    n = 128
    b1 = (ctypes.c_ulonglong * (n * 6))()
    _lib.call("gtk_debug_wn_trace", b1, n)
    wn = np.array(list(b1), dtype=np.int64).reshape(n, 6)
    b2 = (ctypes.c_ulonglong * (n * 5))()
    _lib.call("gtk_debug_conv_trace", b2, n)
    cv = np.array(list(b2), dtype=np.int64).reshape(n, 5)
    wn = wn[wn[:, 0] > wn[:, 0].max() - 200_000]
    cv = cv[cv[:, 0] > cv[:, 0].max() - 200_000]
    end_wn = wn[:, 5].max()
    print(f"wn_layer (last): first CTA past pdl_wait {0:.2f}, median CTA end {(np.median(wn[:,5]) - wn[:,1].min())/1e3:.2f} us, "
          f"last CTA end {(end_wn - wn[:,1].min())/1e3:.2f} us after it")
    print(f"end conv: CTAs launched {(cv[:,0].min() - end_wn)/1e3:+.2f} us relative to the WN kernel's last CTA end (negative = PDL early start)")
    print(f"end conv: first CTA past pdl_wait {(cv[:,1].min() - end_wn)/1e3:+.2f} us, median {(np.median(cv[:,1]) - end_wn)/1e3:+.2f} us")
    print(f"end conv: accumulator complete (median) {(np.median(cv[:,2]) - end_wn)/1e3:+.2f} us, epilogue done {(np.median(cv[:,3]) - end_wn)/1e3:+.2f}, "
          f"stores issued {(np.median(cv[:,4]) - end_wn)/1e3:+.2f} us, last {(cv[:,4].max() - end_wn)/1e3:+.2f}")
