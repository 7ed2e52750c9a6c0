import torch, sys, os
sys.path.insert(0, ".")
from glowtts_b200 import ops
dev = torch.device("cuda", 0)
def run(B, TX, TY):
    g = torch.Generator(device=dev).manual_seed(1)
    value = torch.randn(B, TX, TY, device=dev, generator=g)
    t_xs = torch.full((B,), TX, device=dev, dtype=torch.int32); t_ys = torch.full((B,), TY, device=dev, dtype=torch.int32)
    for _ in range(3): out = ops.mas_path(value, t_xs, t_ys, 0)
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = ops.mas_path(value, t_xs, t_ys, 0); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"B={B} T_x={TX} T_y={TY}: {ts[len(ts)//2]:8.1f} us   {ts[len(ts)//2]*1e3/TY:6.1f} ns/column")
for shape in [(1, 64, 1600), (1, 64, 3200), (1, 128, 3200), (1, 180, 3200), (32, 180, 860)]:
    run(*shape)
