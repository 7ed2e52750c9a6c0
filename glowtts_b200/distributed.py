"""Data-parallel plumbing of the hot path (train_ms_emo_lang_pitch.py:150-170: one process per GPU, DDP).

The hot path shards by utterance: MAS, logp and the decoder forward need no exchange at all; the only collective
is the all-reduce of the decoder's flat gradient buffer (ONE bucket, ops.decoder_backward_flat).  These helpers are
backend-agnostic (NCCL on the GPU box, gloo in the CPU tests) and hold no CUDA code.
"""
import os
from typing import Optional, Tuple

import torch


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of n utterances for `rank` of `world`; sizes differ by at most one and the shards
    tile [0, n) in rank order (DistributedSampler without shuffling or padding)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"glowtts_b200: bad rank/world {rank}/{world}")
    base, rem = divmod(max(n, 0), world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def init_from_env(backend: Optional[str] = None, device: Optional[torch.device] = None) -> Tuple[int, int]:
    """Join the torchrun rendezvous described by RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT (address defaults to
    127.0.0.1: single node).  -> (rank, world).  A world of 1 does not create a process group."""
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = backend or ("nccl" if (device is not None and device.type == "cuda") else "gloo")
        kw = {"device_id": device} if backend == "nccl" and device is not None else {}
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world


def sync_flat_grad(flat: torch.Tensor, group=None) -> torch.Tensor:
    """Average the flat decoder gradient over the ranks of `group`, in place (what DDP does to its buckets;
    here the bucket is the whole decoder, so it is one collective launch per step)."""
    import torch.distributed as dist
    if not dist.is_initialized():
        return flat
    world = dist.get_world_size(group)
    if world > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.div_(world)
    return flat


def max_over_ranks(value: float, device: torch.device) -> float:
    """Device-side timing rule: a multi-GPU step takes as long as its slowest rank."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, device: torch.device) -> float:
    """Whole-job unit count (frames, cells) = sum of the per-rank shard counts."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
