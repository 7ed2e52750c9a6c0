"""MAS kernel timing (CUDA events) for BASELINE configs 1 and 5, every kernel path.
usage: python profiles/mas_bench.py [B5]   (B5 = batch of the cfg-5 run, default 512)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from glowtts_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
PEAK = 6546.6


def run(B, TX, TY, ragged, kernels=(0, 2, 1), reps=10):
    gen = torch.Generator(device=dev).manual_seed(1234)
    value = torch.randn(B, TX, TY, device=dev, generator=gen)
    if ragged:
        t_xs = torch.randint(TX // 2, TX + 1, (B,), device=dev, generator=gen, dtype=torch.int32)
        t_ys = (torch.randint(TY // 4, TY // 2 + 1, (B,), device=dev, generator=gen, dtype=torch.int32) * 2)
    else:
        t_xs = torch.full((B,), TX, device=dev, dtype=torch.int32)
        t_ys = torch.full((B,), TY, device=dev, dtype=torch.int32)
    cells = float((t_xs.long() * t_ys.long()).sum())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ref = None
    for k in kernels:
        for _ in range(2):
            out = ops.mas_path(value, t_xs, t_ys, k)
        ts = []
        for _ in range(reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = ops.mas_path(value, t_xs, t_ys, k)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        us = ts[len(ts) // 2]
        if ref is None:
            ref = out[0]
        else:
            assert torch.equal(ref, out[0])
        gbs = 8 * B * TX * TY / us / 1e3
        print(f"B={B} T_x={TX} T_y={TY} ragged={int(ragged)} kernel={k}: {us:9.1f} us  {cells / us / 1e3:8.2f} Gcells/s (valid)"
              f"  {gbs:7.1f} GB/s padded-plane traffic = {gbs / PEAK:.3f} of HBM peak")


run(32, 180, 860, False)
run(32, 180, 860, True)
b5 = int(sys.argv[1]) if len(sys.argv) > 1 else 512
run(b5, 400, 2000, False, kernels=(0, 2))
run(b5, 400, 2000, True, kernels=(0, 2))
run(256, 256, 1000, False, kernels=(0, 2))
