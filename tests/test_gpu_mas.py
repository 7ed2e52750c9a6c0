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


# ------------------------------------------------------------------------------------------ fused logp + MAS
def _fused_case(B, TX, TY, seed):
    gen = torch.Generator().manual_seed(seed)
    x_m = torch.randn(B, 80, TX, generator=gen) * 0.5
    z = torch.randn(B, 80, TY, generator=gen)
    t_xs = torch.randint(max(1, TX // 2), TX + 1, (B,), generator=gen)
    t_ys = torch.randint(max(TX, TY // 2), TY + 1, (B,), generator=gen)
    t_xs[0], t_ys[0] = TX, TY
    t_ys = torch.maximum(t_ys, t_xs)
    return x_m, z, t_xs.to(torch.int32), t_ys.to(torch.int32)


@pytest.mark.parametrize("B,TX,TY", [(3, 37, 100), (2, 128, 256), (3, 130, 300), (2, 200, 64 * 4), (2, 256, 1000),
                                     (4, 180, 860), (2, 1, 8), (2, 33, 36)])
def test_fused_logp_mas_agrees_with_logp_then_mas(B, TX, TY):
    """gtk_mas_fused_logp (north_star: 'fused variant'): reported as max |delta logp| and path agreement, plus the
    property that makes agreement meaningful -- the fused path is a valid monotonic alignment whose total score under
    the ORACLE logp equals the oracle path's optimum up to fp32 rounding."""
    from glowtts_b200 import monotonic_align, ops
    from oracle import flow as oflow
    x_m, z, t_xs, t_ys = _fused_case(B, TX, TY, 100 + TX)
    assert ops.mas_fused_supported(80, TX, TY, False)
    path, dur, idx, lp = monotonic_align.maximum_path_from_prior(z.cuda(), x_m.cuda(), None, t_xs.cuda(), t_ys.cuda(),
                                                                 return_logp=True, fused=True)
    torch.cuda.synchronize()
    path, dur, idx, lp = path.cpu().numpy(), dur.cpu().numpy(), idx.cpu().numpy(), lp.cpu().numpy()
    want_lp = oflow.logp(x_m, torch.zeros_like(x_m), z).numpy()
    mask = np.zeros((B, TX, TY), np.float32)
    for i in range(B):
        mask[i, : t_xs[i], : t_ys[i]] = 1
    d_logp = np.abs(lp - want_lp * mask).max()
    assert d_logp < 5e-4, d_logp                                    # |logp| ~ 1e2: fp32-level agreement from 3xTF32
    ref = omas.maximum_path_np(want_lp * mask, mask)
    agree, total = 0, 0
    for i in range(B):
        tx, ty = int(t_xs[i]), int(t_ys[i])
        p = path[i]
        assert set(np.unique(p)) <= {0.0, 1.0}
        assert p[tx:].sum() == 0 and p[:, ty:].sum() == 0
        assert (p[:tx, :ty].sum(0) == 1).all()                      # one token per frame
        tok = p[:tx, :ty].argmax(0)
        assert tok[0] == 0 and tok[-1] == tx - 1 and ((np.diff(tok) == 0) | (np.diff(tok) == 1)).all()
        assert (dur[i, :tx] == p[:tx, :ty].sum(1)).all() and (idx[i, :ty] == tok).all() and (idx[i, ty:] == -1).all()
        score = float((want_lp[i].astype(np.float64) * p).sum())
        best = float((want_lp[i].astype(np.float64) * ref[i]).sum())
        assert score >= best - 1e-3 * max(1.0, abs(best)) * 1e-2, (score, best)     # optimal up to rounding (1e-5 rel)
        agree += int((tok == ref[i, :tx, :ty].argmax(0)).sum())
        total += ty
    print(f"fused MAS B={B} T_x={TX} T_y={TY}: max|dlogp|={d_logp:.2e}, path agreement {agree / total:.4f}")
    assert agree / total > 0.97


def test_fused_wrapper_routes_uncovered_shapes_through_logp_then_mas():
    from glowtts_b200 import monotonic_align, ops
    gen = torch.Generator().manual_seed(5)
    B, TX, TY = 2, 300, 640                                          # T_x > 256: not fused
    assert not ops.mas_fused_supported(80, TX, TY, False)
    assert not ops.mas_fused_supported(80, 100, 400, True)           # x_logs given: not fused
    assert not ops.mas_fused_supported(80, 100, 402, False)          # T_y % 4 != 0
    x_m = torch.randn(B, 80, TX, generator=gen).cuda() * 0.5
    z = torch.randn(B, 80, TY, generator=gen).cuda()
    t_xs = torch.tensor([300, 211], dtype=torch.int32).cuda()
    t_ys = torch.tensor([640, 500], dtype=torch.int32).cuda()
    path, dur, idx = monotonic_align.maximum_path_from_prior(z, x_m, None, t_xs, t_ys)
    lp = ops.logp(x_m, None, z)
    want, _, _ = ops.mas_path(lp, t_xs, t_ys)
    assert torch.equal(path, want)
    with pytest.raises(RuntimeError):
        ops.mas_fused_logp(z, x_m, t_xs, t_ys)


@pytest.mark.parametrize("shape", [(32, 180, 860), (3, 33, 65), (2, 600, 700)])
def test_index_only_mode_matches_full_mode(shape):
    """gtk_mas_forward with path = NULL (ops.mas_index): same durations and per-frame index as the full call, on the
    TMA, helper-warp and generic kernels."""
    from glowtts_b200 import ops
    b, tx, ty = shape
    value, mask, t_xs, t_ys = helpers.mas_case(b, tx, ty, 900 + tx, ragged=True, quantised=True)
    v = torch.from_numpy(value).cuda()
    xs, ys = torch.from_numpy(t_xs).cuda(), torch.from_numpy(t_ys).cuda()
    path, dur, idx = ops.mas_path(v, xs, ys)
    dur2, idx2 = ops.mas_index(v, xs, ys)
    assert torch.equal(dur, dur2) and torch.equal(idx, idx2)
    assert torch.equal(path.sum(2).to(torch.int32), dur2)
