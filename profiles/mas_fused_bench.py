import sys, json, torch
sys.path.insert(0, '.')
import bench
print(json.dumps(bench.mas_fused_alone(torch.device('cuda', 0)), indent=1))
