"""Where the stream-K weight-gradient kernel (csrc/wgrad_group.cu) spends its cycles, per CTA of the step's last launch.
    GTK_LIB=profiles/_instr/libglowtts_b200.so python profiles/gg_trace_r02.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from glowtts_b200 import _lib

run = bench.Runner(torch.device("cuda", 0), 1, "bf16", graph=False)
for _ in range(3):
    run.step_resident()
torch.cuda.synchronize()
n = 148
buf = (ctypes.c_longlong * (n * 8))()
_lib.call("gtk_debug_gg_wait", buf, n)
w = np.array(list(buf), dtype=np.int64).reshape(n, 8)
names = ["MMA warp waiting for operand stages", "MMA warp waiting for the drain", "MMA warp total", "k-steps", "segments",
         "entry -> MMA start", "producer waiting for free stages", "drain time (warp 4)"]
print(f"{'per CTA (SM cycles)':40s} {'min':>9s} {'median':>9s} {'max':>9s}")
for i, nme in enumerate(names):
    print(f"{nme:40s} {w[:, i].min():9d} {int(np.median(w[:, i])):9d} {w[:, i].max():9d}")
issue = w[:, 2] - w[:, 0] - w[:, 1]
print("MMA warp issuing (total - waits): median %d cycles = %.1f per k-step" % (np.median(issue), np.median(issue / np.maximum(w[:, 3], 1))))
