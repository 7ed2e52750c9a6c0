"""One train-step of the hot path (bench.py's Runner) for ncu: `python profiles/run_step.py [precision] [n_steps]`."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402

precision = sys.argv[1] if len(sys.argv) > 1 else "bf16"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2
torch.cuda.set_device(0)
run = bench.Runner(torch.device("cuda", 0), 1, precision)
for _ in range(n):
    run.step_resident()
torch.cuda.synchronize()
print("ok")
