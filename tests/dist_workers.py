"""Worker functions of the world_size-2 gloo tests (top-level so torch.multiprocessing.spawn can pickle them)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def _join(rank, world, port):
    from glowtts_b200 import distributed as gd
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    r, w = gd.init_from_env(backend="gloo")
    assert (r, w) == (rank, world) and dist.get_backend() == "gloo"
    return gd


def reductions(rank, world, port, out_dir):
    gd = _join(rank, world, port)
    flat = torch.arange(10, dtype=torch.float32) * (rank + 1)            # rank r holds (r+1) * [0..9]
    gd.sync_flat_grad(flat, dist.group.WORLD)
    mx = gd.max_over_ranks(10.0 + rank, torch.device("cpu"))
    sm = gd.sum_over_ranks(100.0 * (rank + 1), torch.device("cpu"))
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), flat=flat.numpy(), mx=mx, sm=sm)
    dist.barrier()
    dist.destroy_process_group()


def mas_sharded(rank, world, port, out_dir):
    """Each rank aligns its utterance shard with no exchange on the data path; rank 0 gathers for the check."""
    import helpers
    from oracle import mas as omas
    gd = _join(rank, world, port)
    value, mask, _, _ = helpers.mas_case(7, 23, 61, seed=11, ragged=True)       # same seeded batch on every rank
    lo, hi = gd.shard_bounds(value.shape[0], rank, world)
    mine = omas.maximum_path_np(value[lo:hi], mask[lo:hi])
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, mine))
    if rank == 0:
        full = np.zeros_like(value)
        for a, b, p in parts:
            full[a:b] = p
        np.savez(os.path.join(out_dir, "mas.npz"), sharded=full, whole=omas.maximum_path_np(value, mask))
    dist.barrier()
    dist.destroy_process_group()


def flat_grad_data_parallel(rank, world, port, out_dir):
    """The decoder's data-parallel step on the CPU oracle: per-rank shard -> local gradients packed in the flat
    layout the CUDA backward writes (ops.grad_layout) -> ONE all-reduce -> every rank holds the rank-mean."""
    import helpers
    from glowtts_b200 import ops
    from oracle import flow as oflow
    gd = _join(rank, world, port)
    cfg = dict(helpers.TINY_CFG)
    sd = helpers.make_state_dict(cfg, 5)
    names = [k for k in sd if ".wn_pitch." not in k and ".wn_energy." not in k]
    y, mask, _, _ = helpers.flow_inputs(cfg, 4, 24, 3, False)

    def local_flat(lo, hi):
        prm = {k: sd[k].clone().requires_grad_(True) for k in names}
        z, ld = oflow.decoder(prm, y[lo:hi], mask[lo:hi], hidden=cfg["hidden_channels"], dilation_rate=cfg["dilation_rate"],
                              n_blocks=cfg["n_blocks"], n_layers=cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"])
        oflow.mle_loss(z, torch.zeros_like(z), torch.zeros_like(z), ld, mask[lo:hi]).backward()
        plist = [prm[k] for k in names]
        offs, sizes, total = ops.grad_layout(plist)
        flat = torch.zeros(total)
        for p, o, n in zip(plist, offs, sizes):
            flat[o:o + n] = p.grad.reshape(-1)
        return flat

    lo, hi = gd.shard_bounds(y.shape[0], rank, world)
    flat = local_flat(lo, hi)
    gd.sync_flat_grad(flat, dist.group.WORLD)
    if rank == 0:
        want = sum(local_flat(*gd.shard_bounds(y.shape[0], r, world)) for r in range(world)) / world
        np.savez(os.path.join(out_dir, "dp.npz"), got=flat.numpy(), want=want.numpy())
    else:
        np.savez(os.path.join(out_dir, "dp_r1.npz"), got=flat.numpy())
    dist.barrier()
    dist.destroy_process_group()
