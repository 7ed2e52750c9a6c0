"""N > 1 host logic on CPU: world_size-2 gloo runs of the sharding and gradient-sync helpers the GPU path uses
(glowtts_b200/distributed.py), against the CPU oracle.  The CUDA kernels themselves are covered by the -m gpu tests;
here the collective plumbing, shard bounds and "no exchange on the MAS path" are what is checked."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import dist_workers
from glowtts_b200 import distributed as gd


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, tmp_path, world=2):
    mp.spawn(fn, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)


@pytest.mark.parametrize("n,world", [(32, 1), (32, 2), (32, 8), (7, 2), (5, 8), (0, 4), (512, 8)])
def test_shard_bounds_tile_the_batch(n, world):
    spans = [gd.shard_bounds(n, r, world) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    sizes = [hi - lo for lo, hi in spans]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        gd.shard_bounds(n, world, world)


def test_helpers_are_identity_without_a_process_group():
    flat = torch.arange(4.0)
    assert gd.sync_flat_grad(flat.clone()).equal(flat)
    assert gd.max_over_ranks(3.5, torch.device("cpu")) == 3.5
    assert gd.sum_over_ranks(2.0, torch.device("cpu")) == 2.0


def test_gloo_world2_reductions(tmp_path):
    _spawn(dist_workers.reductions, tmp_path)
    r0, r1 = (np.load(tmp_path / f"r{r}.npz") for r in (0, 1))
    want = np.arange(10, dtype=np.float32) * 1.5                         # mean of 1x and 2x
    np.testing.assert_array_equal(r0["flat"], want)
    np.testing.assert_array_equal(r1["flat"], want)
    assert float(r0["mx"]) == float(r1["mx"]) == 11.0                    # slowest rank
    assert float(r0["sm"]) == float(r1["sm"]) == 300.0                   # whole-job unit count


def test_gloo_world2_mas_shards_need_no_exchange(tmp_path):
    _spawn(dist_workers.mas_sharded, tmp_path)
    g = np.load(tmp_path / "mas.npz")
    np.testing.assert_array_equal(g["sharded"], g["whole"])              # bit-exact, utterances are independent


def test_gloo_world2_flat_gradient_sync(tmp_path):
    _spawn(dist_workers.flat_grad_data_parallel, tmp_path)
    g, g1 = np.load(tmp_path / "dp.npz"), np.load(tmp_path / "dp_r1.npz")
    np.testing.assert_allclose(g["got"], g["want"], rtol=1e-6, atol=1e-8)
    np.testing.assert_array_equal(g["got"], g1["got"])                   # both ranks hold the same averaged buffer
    assert np.abs(g["got"]).max() > 0


@pytest.mark.parametrize("nb,segs", [(12, 3), (12, 1), (12, 2), (12, 4), (12, 12), (12, 20), (5, 3), (5, 5), (2, 3), (1, 3)])
def test_backward_range_bounds_tile_the_blocks(nb, segs, monkeypatch):
    """ops.range_bounds: the block ranges of the overlapped data-parallel backward (issued from the last block down) tile
    0..nb-1, none is empty, and the LAST range issued (blocks bounds[0]..bounds[1]-1, the only exposed collective) is the
    smallest; GTK_GRAD_SYNC_LAST overrides its size within what the segment count allows."""
    from glowtts_b200 import ops
    monkeypatch.delenv("GTK_GRAD_SYNC_LAST", raising=False)
    b = ops.range_bounds(nb, segs)
    assert b[0] == 0 and b[-1] == nb and len(b) == min(max(segs, 1), nb) + 1
    sizes = [hi - lo for lo, hi in zip(b, b[1:])]
    assert all(s >= 1 for s in sizes) and sizes[0] == min(sizes)
    if (nb, segs) == (12, 3):
        assert b == [0, 1, 6, 12]                      # 11..6 | 5..1 | 0
    monkeypatch.setenv("GTK_GRAD_SYNC_LAST", "2")
    b2 = ops.range_bounds(nb, segs)
    assert b2[0] == 0 and b2[-1] == nb and all(hi > lo for lo, hi in zip(b2, b2[1:]))
    if len(b2) > 2:
        assert b2[1] == min(2, nb - (len(b2) - 1) + 1)
