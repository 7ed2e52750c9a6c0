"""Pin oracle/flow.py (CPU fp32 restatement) against fixtures produced by the REFERENCE
modules (tests/golden/make_golden.py) and, when /root/reference is present, against the
imported reference modules directly.  CPU only."""
import os

import numpy as np
import pytest
import torch

import helpers
import refimport
from oracle import flow as oflow

TOL = dict(rtol=1e-4, atol=2e-5)

CASES = {
    "flow_tiny": (helpers.TINY_CFG, False),
    "flow_tiny_g": (dict(helpers.TINY_CFG, gin_channels=6), True),
    "flow_tiny_k3_dil2": (dict(helpers.TINY_CFG, kernel_size=3, dilation_rate=2, n_layers=3), False),
    "flow_base": (helpers.BASE_CFG, False),
}


def _kw(cfg):
    return dict(hidden=cfg["hidden_channels"], dilation_rate=cfg["dilation_rate"], n_blocks=cfg["n_blocks"],
                n_layers=cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"])


@pytest.mark.parametrize("name", list(CASES))
def test_decoder_matches_reference_fixture(golden_dir, name):
    cfg, use_g = CASES[name]
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    sd = helpers.make_state_dict(cfg, int(g["sd_seed"]), end_std=float(g["end_std"]))
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    y, mask, _, cond = helpers.flow_inputs(cfg, int(g["b"]), int(g["t"]), int(g["in_seed"]), use_g)
    y = y.clone().requires_grad_(True)
    z, logdet = oflow.decoder(sd, y, mask, g=cond, **_kw(cfg))
    np.testing.assert_allclose(z.detach().numpy(), g["z"], **TOL)
    np.testing.assert_allclose(logdet.detach().numpy(), g["logdet"], rtol=1e-4, atol=1e-3)
    loss = oflow.mle_loss(z, torch.zeros_like(z), torch.zeros_like(z), logdet, mask)
    assert abs(loss.item() - float(g["loss"])) < 1e-4 * max(1.0, abs(float(g["loss"])))
    loss.backward()
    np.testing.assert_allclose(y.grad.numpy(), g["dy"], rtol=1e-3, atol=1e-6)
    n_checked = 0
    for key in g.files:
        if not key.startswith("grad:"):
            continue
        got = sd[key[5:]].grad.numpy()
        want = g[key]
        if got.size != want.size:
            got = got.reshape(-1)[:: max(1, got.size // 2048)]
        np.testing.assert_allclose(got.reshape(want.shape), want, rtol=2e-3, atol=2e-6)
        n_checked += 1
    assert n_checked > 5
    with torch.no_grad():
        y_rev, none = oflow.decoder(sd, z.detach(), mask, g=cond, reverse=True, **_kw(cfg))
    assert none is None
    np.testing.assert_allclose(y_rev.numpy(), g["y_rev"], rtol=1e-4, atol=5e-5)
    np.testing.assert_allclose(y_rev.numpy(), g["y_rev_stored"], rtol=1e-4, atol=5e-5)


PROSODY = {
    "flow_tiny_pe": dict(helpers.TINY_CFG, gin_channels=6),
    "flow_tiny_p": helpers.TINY_CFG,
}


@pytest.mark.parametrize("name", list(PROSODY))
def test_prosody_decoder_matches_reference_fixture(golden_dir, name):
    """3-WN coupling blocks of the live fork (wn -> wn_energy -> wn_pitch, attentions.py:155-160; modules.WNP)."""
    cfg = PROSODY[name]
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    sd = {k: v.clone().requires_grad_(True) for k, v in helpers.make_state_dict(cfg, int(g["sd_seed"])).items()}
    y, mask, _, cond = helpers.flow_inputs(cfg, int(g["b"]), int(g["t"]), int(g["in_seed"]), bool(g["use_g"]))
    pitch, energy = helpers.prosody_inputs(int(g["b"]), int(g["t"]), int(g["in_seed"]), bool(g["use_pitch"]),
                                           bool(g["use_energy"]))
    y = y.clone().requires_grad_(True)
    pitch = pitch.clone().requires_grad_(True)
    z, logdet = oflow.decoder(sd, y, mask, g=cond, pitch=pitch, energy=energy, **_kw(cfg))
    np.testing.assert_allclose(z.detach().numpy(), g["z"], **TOL)
    np.testing.assert_allclose(logdet.detach().numpy(), g["logdet"], rtol=1e-4, atol=1e-3)
    loss = oflow.mle_loss(z, torch.zeros_like(z), torch.zeros_like(z), logdet, mask)
    loss.backward()
    np.testing.assert_allclose(y.grad.numpy(), g["dy"], rtol=1e-3, atol=1e-6)
    np.testing.assert_allclose(pitch.grad.numpy(), g["dpitch"], rtol=1e-3, atol=1e-6)
    keys = [k for k in g.files if k.startswith("grad:")]
    assert any(".wn_pitch.cond_layer1." in k for k in keys)
    for key in keys:
        np.testing.assert_allclose(sd[key[5:]].grad.numpy(), g[key], rtol=2e-3, atol=2e-6, err_msg=key)
    with torch.no_grad():
        y_rev, _ = oflow.decoder(sd, z.detach(), mask, g=cond, pitch=pitch, energy=energy, reverse=True, **_kw(cfg))
    np.testing.assert_allclose(y_rev.numpy(), g["y_rev"], rtol=1e-4, atol=5e-5)
    np.testing.assert_allclose(y_rev.numpy(), g["y_rev_stored"], rtol=1e-4, atol=5e-5)


def test_modules_match_reference_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "modules_small.npz"))
    x = torch.from_numpy(g["x"])
    lengths = torch.from_numpy(g["lengths"])
    mask = oflow.sequence_mask(lengths, x.shape[2]).float().unsqueeze(1)
    z, ld = oflow.actnorm(x, mask, torch.from_numpy(g["an_logs"]), torch.from_numpy(g["an_bias"]))
    np.testing.assert_allclose(z.numpy(), g["an_z"], **TOL)
    np.testing.assert_allclose(ld.numpy(), g["an_logdet"], **TOL)
    zr, _ = oflow.actnorm(x, mask, torch.from_numpy(g["an_logs"]), torch.from_numpy(g["an_bias"]), reverse=True)
    np.testing.assert_allclose(zr.numpy(), g["an_zrev"], **TOL)
    logs, bias = oflow.actnorm_ddi(x, mask)
    np.testing.assert_allclose(logs.numpy(), g["ddi_logs"], **TOL)
    np.testing.assert_allclose(bias.numpy(), g["ddi_bias"], **TOL)
    w = torch.from_numpy(g["ic_w"])
    z, ld = oflow.invconvnear(x, mask, w)
    np.testing.assert_allclose(z.numpy(), g["ic_z"], **TOL)
    np.testing.assert_allclose(ld.numpy(), g["ic_logdet"], **TOL)
    zr, _ = oflow.invconvnear(x, mask, w, reverse=True)
    np.testing.assert_allclose(zr.numpy(), g["ic_zrev"], **TOL)
    out = oflow.fused_add_tanh_sigmoid_multiply(torch.from_numpy(g["gate_a"]), torch.from_numpy(g["gate_b"]), 16)
    np.testing.assert_allclose(out.numpy(), g["gate_out"], **TOL)
    xs, ms = oflow.squeeze(x, mask, 2)
    assert np.array_equal(xs.numpy(), g["sq_x"]) and np.array_equal(ms.numpy(), g["sq_mask"])
    xu, mu = oflow.unsqueeze(xs, ms, 2)
    assert np.array_equal(xu.numpy(), g["unsq_x"]) and np.array_equal(mu.numpy(), g["unsq_mask"])
    gp = oflow.generate_path(torch.from_numpy(g["gp_dur"]), torch.from_numpy(g["gp_mask"]))
    assert np.array_equal(gp.numpy(), g["gp_path"])


def test_logp_matches_reference_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "logp_small.npz"))
    lp = oflow.logp(torch.from_numpy(g["x_m"]), torch.from_numpy(g["x_logs"]), torch.from_numpy(g["z"]))
    np.testing.assert_allclose(lp.numpy(), g["logp"], rtol=1e-5, atol=1e-4)


def test_expand_prior_is_gather():
    value, mask, t_xs, t_ys = helpers.mas_case(2, 7, 15, 3)
    from oracle import mas as omas
    attn = torch.from_numpy(omas.maximum_path_np(value, mask).astype(np.float32))
    x_m = torch.randn(2, 5, 7)
    z_m, _ = oflow.expand_prior(attn, x_m, x_m)
    idx = attn.argmax(1)
    for b in range(2):
        want = x_m[b][:, idx[b]] * (attn[b].sum(0) > 0)
        assert torch.allclose(z_m[b], want)


@pytest.mark.skipif(not refimport.available(), reason="/root/reference absent")
def test_coupling_and_wn_against_live_reference():
    ref = refimport.load()
    torch.manual_seed(5)
    cb = ref.attentions.CouplingBlock(16, 12, 5, 2, 3, gin_channels=4, p_dropout=0.0).eval()
    cb.end.weight.data.normal_(0, 0.1)
    cb.end.bias.data.normal_(0, 0.1)
    sd = {"cb." + k: v for k, v in cb.state_dict().items()}
    x = torch.randn(2, 16, 9)
    mask = oflow.sequence_mask(torch.tensor([9, 5]), 9).float().unsqueeze(1)
    g = torch.randn(2, 4, 1)
    pitch = torch.randn(2, 1, 18)
    for kwargs in (dict(), dict(pitch=pitch), dict(pitch=pitch, energy=pitch * 0.5)):
        z_ref, ld_ref = cb(x, mask, g=g, **kwargs)
        z, ld = oflow.coupling(sd, "cb", x, mask, hidden=12, dilation_rate=2, n_layers=3, g=g, **kwargs)
        assert torch.allclose(z, z_ref, atol=1e-5) and torch.allclose(ld, ld_ref, atol=1e-4)
        zr_ref, _ = cb(z_ref, mask, g=g, reverse=True, **kwargs)
        zr, _ = oflow.coupling(sd, "cb", z, mask, hidden=12, dilation_rate=2, n_layers=3, g=g, reverse=True, **kwargs)
        assert torch.allclose(zr, zr_ref, atol=1e-5)


@pytest.mark.parametrize("norm_type,clip", [(2, None), (2, 0.05), (1, 0.1), (3, 0.02)])
def test_clip_grad_value_restatement_matches_reference(norm_type, clip):
    """oracle.flow.clip_grad_value_ vs the reference's commons.clip_grad_value_ (commons.py:320-336)."""
    ref = refimport.load()
    if ref is None:
        pytest.skip("/root/reference not present")
    gen = torch.Generator().manual_seed(3)
    grads = [torch.randn(n, generator=gen) * 0.1 for n in (5, 1000, 33)]
    params = []
    for g in grads:
        p = torch.nn.Parameter(torch.zeros_like(g))
        p.grad = g.clone()
        params.append(p)
    want = ref.commons.clip_grad_value_(params, clip, norm_type)
    mine = [g.clone() for g in grads]
    got = oflow.clip_grad_value_(mine, clip, norm_type)
    assert abs(got - want) <= 1e-6 * max(1.0, abs(want))
    for p, m in zip(params, mine):
        assert torch.equal(p.grad, m)


# ---------------------------------------------------------------------------------- round-2 fixtures
ROUND2 = {
    "flow_tiny_sig": (helpers.TINY_CFG, True),
    "flow_h64": (dict(in_channels=16, hidden_channels=64, kernel_size=5, dilation_rate=1, n_blocks=3, n_layers=3,
                      n_split=4, n_sqz=2, gin_channels=0), False),
    "flow_h64_g_sig": (dict(in_channels=16, hidden_channels=64, kernel_size=3, dilation_rate=2, n_blocks=2, n_layers=3,
                            n_split=4, n_sqz=2, gin_channels=8), True),
}


@pytest.mark.parametrize("name", list(ROUND2))
def test_round2_decoder_fixtures_incl_sigmoid_scale(golden_dir, name):
    """sigmoid_scale=True (attentions.py:175-176) and the hidden-64 shapes the tcgen05 path can run."""
    cfg, sig = ROUND2[name]
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    assert bool(g["sigmoid_scale"]) == sig
    sd = {k: v.clone().requires_grad_(True) for k, v in helpers.make_state_dict(cfg, int(g["sd_seed"])).items()}
    y, mask, _, cond = helpers.flow_inputs(cfg, int(g["b"]), int(g["t"]), int(g["in_seed"]), bool(g["use_g"]))
    y = y.clone().requires_grad_(True)
    if cond is not None:
        cond = cond.clone().requires_grad_(True)
    z, logdet = oflow.decoder(sd, y, mask, g=cond, sigmoid_scale=sig, **_kw(cfg))
    np.testing.assert_allclose(z.detach().numpy(), g["z"], **TOL)
    np.testing.assert_allclose(logdet.detach().numpy(), g["logdet"], rtol=1e-4, atol=1e-3)
    loss = oflow.mle_loss(z, torch.zeros_like(z), torch.zeros_like(z), logdet, mask)
    assert abs(loss.item() - float(g["loss"])) < 1e-4 * max(1.0, abs(float(g["loss"])))
    loss.backward()
    np.testing.assert_allclose(y.grad.numpy(), g["dy"], rtol=1e-3, atol=1e-6)
    if cond is not None:
        np.testing.assert_allclose(cond.grad.numpy(), g["dg"], rtol=1e-3, atol=1e-6)
    for key in [k for k in g.files if k.startswith("grad:")]:
        got, want = sd[key[5:]].grad.numpy(), g[key]
        if got.size != want.size:
            got = got.reshape(-1)[:: max(1, got.size // 2048)]
        np.testing.assert_allclose(got.reshape(want.shape), want, rtol=2e-3, atol=2e-6, err_msg=key)
    with torch.no_grad():
        y_rev, _ = oflow.decoder(sd, z.detach(), mask, g=cond, sigmoid_scale=sig, reverse=True, **_kw(cfg))
    np.testing.assert_allclose(y_rev.numpy(), g["y_rev"], rtol=1e-4, atol=5e-5)


def test_modules_without_mask_and_no_jacobian(golden_dir):
    """x_mask=None (modules.py:585-586, 639-641; attentions.py:134-135) and InvConvNear(no_jacobian=True)."""
    g = np.load(os.path.join(golden_dir, "modules_nomask.npz"))
    x = torch.from_numpy(g["x"])
    ones = torch.ones(x.shape[0], 1, x.shape[2])
    z, ld = oflow.actnorm(x, ones, torch.from_numpy(g["an_logs"]), torch.from_numpy(g["an_bias"]))
    np.testing.assert_allclose(z.numpy(), g["an_z"], **TOL)
    np.testing.assert_allclose(ld.numpy(), g["an_logdet"], **TOL)
    w = torch.from_numpy(g["ic_w"])
    z, ld = oflow.invconvnear(x, ones, w)
    np.testing.assert_allclose(z.numpy(), g["ic_z"], **TOL)
    np.testing.assert_allclose(ld.numpy(), g["ic_logdet"], **TOL)
    mask = oflow.sequence_mask(torch.from_numpy(g["nj_lengths"]), x.shape[2]).float().unsqueeze(1)
    z, _ = oflow.invconvnear(x, mask, w)
    np.testing.assert_allclose(z.numpy(), g["nj_z"], **TOL)
    cfg = dict(helpers.TINY_CFG, n_blocks=1)
    sd = helpers.make_state_dict(cfg, 44)
    for tag, sig in (("cb", False), ("cbs", True)):
        z, ld = oflow.coupling(sd, "flows.2", x, ones, hidden=16, dilation_rate=1, n_layers=2, sigmoid_scale=sig)
        np.testing.assert_allclose(z.numpy(), g[tag + "_z"], **TOL)
        np.testing.assert_allclose(ld.numpy(), g[tag + "_logdet"], rtol=1e-4, atol=1e-4)
        zr, _ = oflow.coupling(sd, "flows.2", x, ones, hidden=16, dilation_rate=1, n_layers=2, sigmoid_scale=sig, reverse=True)
        np.testing.assert_allclose(zr.numpy(), g[tag + "_zrev"], **TOL)
