"""Host enqueue time vs device time of one step (is the step launch-bound?)."""
import os, sys, time
sys.path.insert(0, os.getcwd())
import torch, bench
run = bench.Runner(torch.device("cuda", 0), 1, "bf16")
for _ in range(5): run.step_resident()
torch.cuda.synchronize()
N = 20
t0 = time.perf_counter()
for _ in range(N): run.step_resident()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host enqueue {1e3*(t1-t0)/N:.2f} ms/step, wall incl. drain {1e3*(t2-t0)/N:.2f} ms/step")
# host-only cost of the fused decoder calls: time one step with the GPU idle at start, section by section
from glowtts_b200.hot_path import train_hot_path
y, y_len, x_m, x_len = run.dev
def sect(name, fn, n=20):
    torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(n): r = fn()
    h=time.perf_counter()-t; torch.cuda.synchronize(); w=time.perf_counter()-t
    print(f"{name:28s} host {1e3*h/n:6.2f} ms  wall {1e3*w/n:6.2f} ms")
    return r
out = sect("forward (hot path)", lambda: train_hot_path(run.dec, y, y_len, x_m, None, x_len))
def fb():
    run.opt.zero_grad(set_to_none=True)
    o = train_hot_path(run.dec, y, y_len, x_m, None, x_len); o.loss.backward(); return o
sect("forward+backward", fb)
sect("optimizer", lambda: run.opt.step())
