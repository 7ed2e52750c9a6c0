import sys, os
sys.path.insert(0, os.getcwd())
import torch, bench
from torch.profiler import profile, ProfilerActivity
run = bench.Runner(torch.device("cuda", 0), 1, "bf16")
for _ in range(3): run.step_resident()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    run.step_resident()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60))
