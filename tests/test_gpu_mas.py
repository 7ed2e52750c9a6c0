"""GPU parity of monotonic_align.maximum_path (csrc/mas.cu through the C ABI) against the
committed reference fixtures and the CPU oracle: BIT-EXACT (integer path), incl. ties, ragged
lengths, t_x == t_y, t_x == 1, empty utterances, both kernel paths, and size-independent
properties at BASELINE.json's full sizes."""
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import mas as omas

pytestmark = pytest.mark.gpu

GOLDEN = ["mas_small_ragged", "mas_ties", "mas_square", "mas_tx1", "mas_medium", "mas_wide"]


def _gpu_path(value, mask, force_generic=0):
    from glowtts_b200 import ops
    v = torch.from_numpy(value * mask).cuda()
    m = torch.from_numpy(mask).cuda()
    t_xs, t_ys = ops.mas_lengths(m)
    path, dur, idx = ops.mas_path(v, t_xs, t_ys, int(force_generic))
    torch.cuda.synchronize()
    return path.cpu().numpy(), dur.cpu().numpy(), idx.cpu().numpy(), t_xs.cpu().numpy(), t_ys.cpu().numpy()


@pytest.mark.parametrize("generic", [0, 1, 2])     # TMA ring / generic kernel / helper-warp fast path
@pytest.mark.parametrize("name", GOLDEN)
def test_matches_reference_fixture(golden_dir, name, generic):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    b, tx, ty, seed, ragged, quant = (int(v) for v in g["case"])
    value, mask, t_xs, t_ys = helpers.mas_case(b, tx, ty, seed, bool(ragged), bool(quant))
    path, dur, idx, gx, gy = _gpu_path(value, mask, generic)
    assert (gx == t_xs).all() and (gy == t_ys).all()
    assert set(np.unique(path)) <= {0.0, 1.0}
    assert (helpers.path_to_index(path) == g["index"]).all()
    assert (path.sum((1, 2)) == g["path_sum"]).all()
    assert (dur == path.sum(2)).all()
    want_idx = np.where(path.sum(1) > 0, path.argmax(1), -1)
    assert (idx == want_idx).all()


def test_drop_in_wrapper_matches_reference_wrapper_semantics():
    from glowtts_b200 import monotonic_align
    value, mask, _, _ = helpers.mas_case(3, 33, 77, 41)
    want = omas.maximum_path_np(value, mask)
    for dtype in (torch.float32, torch.float16, torch.bfloat16):
        v = torch.from_numpy(value).cuda().to(dtype)
        m = torch.from_numpy(mask).cuda().to(dtype)
        got = monotonic_align.maximum_path(v, m)
        assert got.dtype == dtype and got.device == v.device and got.shape == v.shape
        ref = omas.maximum_path_np((v * m).float().cpu().numpy(), mask)   # same rounding of the inputs
        assert (got.float().cpu().numpy() == ref).all()
    assert (monotonic_align.maximum_path(torch.from_numpy(value).cuda(), torch.from_numpy(mask).cuda())
            .cpu().numpy() == want).all()


@pytest.mark.parametrize("shape", [(32, 180, 860), (5, 31, 64), (3, 33, 65), (2, 64, 64), (4, 129, 400),
                                   (2, 400, 2000), (3, 512, 1001), (2, 7, 2048), (1, 1, 1), (2, 300, 301),
                                   (3, 128, 400), (2, 256, 512), (2, 512, 1000), (3, 97, 100), (2, 333, 1500)])
@pytest.mark.parametrize("quant", [False, True])
def test_matches_oracle(shape, quant):
    b, tx, ty = shape
    value, mask, _, _ = helpers.mas_case(b, tx, ty, 1000 + tx + ty, ragged=True, quantised=quant)
    want = omas.maximum_path_np(value, mask, impl="c_port_omp")
    for generic in (0, 1, 2):
        path, dur, idx, _, _ = _gpu_path(value, mask, generic)
        assert (path == want).all(), f"generic={generic}"
        assert (dur == want.sum(2)).all()


def test_large_tx_uses_generic_path():
    value, mask, _, _ = helpers.mas_case(2, 600, 1200, 77, ragged=True)
    want = omas.maximum_path_np(value, mask, impl="c_port_omp")
    path, _, _, _, _ = _gpu_path(value, mask)
    assert (path == want).all()


def test_unaligned_rows_and_odd_ty():
    # T_y % 4 != 0 exercises the scalar loader
    for ty in (61, 62, 63, 127):
        value, mask, _, _ = helpers.mas_case(3, 20, ty, ty, ragged=True, quantised=True)
        want = omas.maximum_path_np(value, mask)
        path, _, _, _, _ = _gpu_path(value, mask)
        assert (path == want).all(), ty


def test_empty_and_degenerate_utterances():
    value = np.ones((3, 6, 10), np.float32)
    mask = np.zeros((3, 6, 10), np.float32)
    mask[1, :2, :4] = 1
    mask[2, :6, :6] = 1
    want = omas.maximum_path_np(value, mask)
    path, dur, idx, _, _ = _gpu_path(value, mask)
    assert (path == want).all()
    assert path[0].sum() == 0 and (idx[0] == -1).all() and (dur[0] == 0).all()
    # zero-sized batch / dims do not crash
    from glowtts_b200 import ops
    z = torch.zeros(0, 4, 5, device="cuda")
    p, d, i = ops.mas_path(z, torch.zeros(0, dtype=torch.int32, device="cuda"), torch.zeros(0, dtype=torch.int32, device="cuda"))
    assert p.shape == (0, 4, 5)


def test_value_is_not_modified():
    value, mask, _, _ = helpers.mas_case(2, 17, 40, 5)
    from glowtts_b200 import monotonic_align
    v = torch.from_numpy(value).cuda()
    keep = v.clone()
    monotonic_align.maximum_path(v, torch.from_numpy(mask).cuda())
    assert torch.equal(v, keep)


def _check_properties(path, idx, dur, t_xs, t_ys):
    b, tx, ty = path.shape
    cols = path.sum(1)
    ar_y = torch.arange(ty, device=path.device)[None, :]
    ar_x = torch.arange(tx, device=path.device)[None, :]
    in_y = ar_y < t_ys[:, None]
    assert torch.equal(cols, in_y.float())                          # one token per frame, none in the padding
    assert torch.equal(dur.float(), path.sum(2))
    assert ((dur > 0) == (ar_x < t_xs[:, None])).all()              # surjective onto [0, t_x)
    step = idx[:, 1:] - idx[:, :-1]
    inside = in_y[:, 1:]
    assert (((step == 0) | (step == 1)) | ~inside).all()            # monotone without skips
    assert (idx[:, 0] == 0).all()
    last = idx.gather(1, (t_ys.long() - 1).clamp(min=0)[:, None])[:, 0]
    assert (last == t_xs - 1).all()


def test_full_size_properties_and_sampled_parity():
    """BASELINE cfg 5 shape (T_x=400, T_y=2000) at B=64 (B=512 is 3.3 GB x2: covered by bench), full + ragged."""
    from glowtts_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(1234)
    B, TX, TY = 64, 400, 2000
    value = torch.randn(B, TX, TY, device="cuda", generator=gen)
    t_xs = torch.randint(200, TX + 1, (B,), device="cuda", generator=gen, dtype=torch.int32)
    t_ys = torch.randint(1000, TY + 1, (B,), device="cuda", generator=gen, dtype=torch.int32)
    t_xs[0], t_ys[0] = TX, TY
    mask = ((torch.arange(TX, device="cuda")[None, :, None] < t_xs[:, None, None]) &
            (torch.arange(TY, device="cuda")[None, None, :] < t_ys[:, None, None])).float()
    v = value * mask
    path, dur, idx = ops.mas_path(v, t_xs, t_ys)
    _check_properties(path, idx, dur, t_xs, t_ys)
    # bit-exact parity on a sample of the utterances
    sel = [0, 1, 17, 63]
    want = omas.maximum_path_np(v[sel].cpu().numpy(), mask[sel].cpu().numpy(), impl="c_port_omp")
    assert (path[sel].cpu().numpy() == want).all()
    # generic path agrees with the fast path everywhere
    for kernel in (1, 2):
        path2, dur2, idx2 = ops.mas_path(v, t_xs, t_ys, kernel)
        assert torch.equal(path, path2) and torch.equal(dur, dur2) and torch.equal(idx, idx2)


def test_duration_roundtrip_through_generate_path():
    """commons.generate_path(attn.sum(2)) reproduces attn (models.py:1085 + commons.py:127-143)."""
    from glowtts_b200 import commons, ops
    value, mask, _, _ = helpers.mas_case(4, 50, 120, 9)
    v, m = torch.from_numpy(value * mask).cuda(), torch.from_numpy(mask).cuda()
    t_xs, t_ys = ops.mas_lengths(m)
    path, dur, _ = ops.mas_path(v, t_xs, t_ys)
    again = commons.generate_path(dur.float(), m)
    assert torch.equal(again, path)
