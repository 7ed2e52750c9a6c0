"""torchrun -N 2 check: the data-parallel backward with deferred range joins (gtk_decoder_backward_range_deferred +
gtk_decoder_backward_join, the default) gives the gradients of the joined ranges (GTK_RANGE_DEFER=0) and of a
single-call backward followed by one all-reduce.  python -m torch.distributed.run --nproc-per-node 2 profiles/ddp_defer_check.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import helpers  # noqa: E402
from glowtts_b200 import models, ops  # noqa: E402

rank = int(os.environ["RANK"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
cfg = dict(in_channels=80, hidden_channels=192, kernel_size=5, dilation_rate=1, n_blocks=12, n_layers=4, n_split=4, n_sqz=2,
           gin_channels=0)
sd = helpers.make_state_dict(cfg, 5)
dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"], cfg["n_blocks"],
                             cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"], precision="bf16")
dec.load_state_dict(sd, strict=True)
dec = dec.cuda().train()
gen = torch.Generator().manual_seed(100 + rank)
lens = torch.tensor([860, 430, 612, 700, 256, 128, 800, 344])
B, T = len(lens), int(lens.max())
mask = (torch.arange(T)[None, :] < lens[:, None]).float().cuda()
y = (torch.randn(B, 80, T, generator=gen).cuda() * mask[:, None, :])
params, folded = dec._params()
flat, offsets = dec._ensure_flat(params)
c = dec._cfg(folded)
out = {}
with torch.no_grad():
    z, logdet, ws = ops.decoder_forward_flat(y, mask, flat, offsets, None, c, True, None)
    dz = torch.randn(z.shape, generator=gen).cuda() * mask[:, None, :z.shape[2]]
    dld = torch.randn(logdet.shape, generator=gen).cuda()
    dx0, df0, *_ = ops.decoder_backward_flat(dz, dld, mask, flat, offsets, None, ws, c, T, None)
    dist.all_reduce(df0, op=dist.ReduceOp.AVG)
    for mode in ("1", "0", "1"):
        os.environ["GTK_RANGE_DEFER"] = mode
        dx1, df1, *_ = ops.decoder_backward_flat_overlapped(dz, dld, mask, flat, offsets, None, ws, c, T, None, dist.group.WORLD, 3)
        torch.cuda.synchronize()
        rel = float((df1.double() - df0.double()).norm() / df0.double().norm())
        relx = float((dx1.double() - dx0.double()).norm() / dx0.double().norm())
        assert rel < 1e-5 and relx < 1e-6, (mode, rel, relx)
        out[mode] = (rel, relx)
if rank == 0:
    print("ddp_defer_check ok", out)
dist.barrier()
os._exit(0)
