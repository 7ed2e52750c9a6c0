"""GPU parity of the flow decoder path (fp32 CUDA path through the C ABI) against the committed
fixtures made by the REFERENCE modules and against the CPU oracle (oracle/flow.py).
Tolerance (BASELINE.json north_star): fp32, 1e-4 relative (absolute floors stated per check)."""
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import flow as oflow

pytestmark = pytest.mark.gpu

TOL = dict(rtol=1e-4, atol=2e-5)
CASES = {
    "flow_tiny": (helpers.TINY_CFG, False),
    "flow_tiny_g": (dict(helpers.TINY_CFG, gin_channels=6), True),
    "flow_tiny_k3_dil2": (dict(helpers.TINY_CFG, kernel_size=3, dilation_rate=2, n_layers=3), False),
    "flow_base": (helpers.BASE_CFG, False),
}


def _decoder(cfg, sd, precision="fp32"):
    from glowtts_b200 import models
    dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                 cfg["n_blocks"], cfg["n_layers"], p_dropout=0.0, n_split=cfg["n_split"],
                                 n_sqz=cfg["n_sqz"], gin_channels=cfg["gin_channels"], precision=precision)
    dec.load_state_dict(sd, strict=True)      # every reference key, incl. the fork's wn_pitch / wn_energy
    return dec.cuda()


def _kw(cfg):
    return dict(hidden=cfg["hidden_channels"], dilation_rate=cfg["dilation_rate"], n_blocks=cfg["n_blocks"],
                n_layers=cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"])


@pytest.mark.parametrize("name", list(CASES))
def test_decoder_matches_reference_fixture(golden_dir, name):
    from glowtts_b200 import commons
    cfg, use_g = CASES[name]
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    sd = helpers.make_state_dict(cfg, int(g["sd_seed"]), end_std=float(g["end_std"]))
    dec = _decoder(cfg, sd)
    y, mask, _, cond = helpers.flow_inputs(cfg, int(g["b"]), int(g["t"]), int(g["in_seed"]), use_g)
    y = y.cuda().requires_grad_(True)
    mask = mask.cuda()
    cond = cond.cuda() if cond is not None else None
    z, logdet = dec(y, mask, g=cond)
    np.testing.assert_allclose(z.detach().cpu().numpy(), g["z"], **TOL)
    np.testing.assert_allclose(logdet.detach().cpu().numpy(), g["logdet"], rtol=1e-4, atol=1e-3)
    loss = commons.mle_loss(z, None, None, logdet, mask)
    assert abs(loss.item() - float(g["loss"])) < 1e-4 * max(1.0, abs(float(g["loss"])))
    loss.backward()
    np.testing.assert_allclose(y.grad.cpu().numpy(), g["dy"], rtol=1e-3, atol=1e-6)
    params = dict(dec.named_parameters())
    n_checked = 0
    for key in g.files:
        if not key.startswith("grad:"):
            continue
        got = params[key[5:]].grad.cpu().numpy()
        want = g[key]
        if got.size != want.size:
            got = got.reshape(-1)[:: max(1, got.size // 2048)]
        np.testing.assert_allclose(got.reshape(want.shape), want, rtol=2e-3, atol=2e-6, err_msg=key)
        n_checked += 1
    assert n_checked > 5
    y_rev, none = dec(z.detach(), mask, g=cond, reverse=True)
    assert none is None
    np.testing.assert_allclose(y_rev.cpu().numpy(), g["y_rev"], rtol=1e-4, atol=5e-5)
    dec.store_inverse()
    y_rev2, _ = dec(z.detach(), mask, g=cond, reverse=True)
    np.testing.assert_allclose(y_rev2.cpu().numpy(), g["y_rev_stored"], rtol=1e-4, atol=5e-5)


PROSODY = {
    "flow_tiny_pe": dict(helpers.TINY_CFG, gin_channels=6),
    "flow_tiny_p": helpers.TINY_CFG,
}


@pytest.mark.parametrize("name", list(PROSODY))
def test_prosody_decoder_matches_reference_fixture(golden_dir, name):
    """The live fork's 3-WN coupling blocks (wn -> wn_energy -> wn_pitch, attentions.py:155-160; modules.WNP
    modules.py:272-362) against fixtures made by the reference modules: z, logdet, loss, every gradient
    (incl. cond_layer1 and the pitch track itself), reverse before and after store_inverse."""
    from glowtts_b200 import commons
    cfg = PROSODY[name]
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    dec = _decoder(cfg, helpers.make_state_dict(cfg, int(g["sd_seed"])))
    y, mask, _, cond = helpers.flow_inputs(cfg, int(g["b"]), int(g["t"]), int(g["in_seed"]), bool(g["use_g"]))
    pitch, energy = helpers.prosody_inputs(int(g["b"]), int(g["t"]), int(g["in_seed"]), bool(g["use_pitch"]),
                                           bool(g["use_energy"]))
    y = y.cuda().requires_grad_(True)
    mask = mask.cuda()
    cond = cond.cuda() if cond is not None else None
    pitch = pitch.cuda().requires_grad_(True)
    energy = energy.cuda() if energy is not None else None
    z, logdet = dec(y, mask, g=cond, pitch=pitch, energy=energy)
    np.testing.assert_allclose(z.detach().cpu().numpy(), g["z"], **TOL)
    np.testing.assert_allclose(logdet.detach().cpu().numpy(), g["logdet"], rtol=1e-4, atol=1e-3)
    loss = commons.mle_loss(z, None, None, logdet, mask)
    assert abs(loss.item() - float(g["loss"])) < 1e-4 * max(1.0, abs(float(g["loss"])))
    loss.backward()
    np.testing.assert_allclose(y.grad.cpu().numpy(), g["dy"], rtol=1e-3, atol=1e-6)
    np.testing.assert_allclose(pitch.grad.cpu().numpy(), g["dpitch"], rtol=1e-3, atol=1e-6)
    params = dict(dec.named_parameters())
    keys = [k for k in g.files if k.startswith("grad:")]
    assert any(".wn_pitch.cond_layer1." in k for k in keys)
    for key in keys:
        np.testing.assert_allclose(params[key[5:]].grad.cpu().numpy(), g[key], rtol=2e-3, atol=2e-6, err_msg=key)
    with_grad = {"grad:" + k for k, p in params.items() if p.grad is not None}
    assert with_grad == set(keys)          # an unconditioned WNP is the identity: its parameters get no gradient
    with torch.no_grad():
        y_rev, none = dec(z.detach(), mask, g=cond, pitch=pitch.detach(), energy=energy, reverse=True)
    assert none is None
    np.testing.assert_allclose(y_rev.cpu().numpy(), g["y_rev"], rtol=1e-4, atol=5e-5)
    dec.store_inverse()
    with torch.no_grad():
        y_rev2, _ = dec(z.detach(), mask, g=cond, pitch=pitch.detach(), energy=energy, reverse=True)
    np.testing.assert_allclose(y_rev2.cpu().numpy(), g["y_rev_stored"], rtol=1e-4, atol=5e-5)


@pytest.mark.parametrize("b,t,use_g", [(3, 50, False), (2, 97, True), (1, 2, False), (4, 33, False)])
def test_decoder_matches_oracle_random(b, t, use_g):
    """Seeded random shapes incl. odd T (last frame dropped by squeeze) and ragged lengths."""
    cfg = dict(helpers.TINY_CFG, n_blocks=3, n_layers=3, hidden_channels=24, gin_channels=5 if use_g else 0)
    sd = helpers.make_state_dict(cfg, 900 + t)
    dec = _decoder(cfg, sd)
    y, mask, _, cond = helpers.flow_inputs(cfg, b, t, 77 + t, use_g)
    z_ref, ld_ref = oflow.decoder({k: v for k, v in sd.items()}, y, mask, g=cond, **_kw(cfg))
    z, ld = dec(y.cuda(), mask.cuda(), g=None if cond is None else cond.cuda())
    np.testing.assert_allclose(z.detach().cpu().numpy(), z_ref.numpy(), **TOL)
    np.testing.assert_allclose(ld.detach().cpu().numpy(), ld_ref.numpy(), rtol=1e-4, atol=1e-3)
    # invertibility: decoder^-1(decoder(y)) == y * mask on the frames squeeze keeps
    y_back, _ = dec(z.detach(), mask.cuda(), g=None if cond is None else cond.cuda(), reverse=True)
    te = (t // 2) * 2
    ms = mask[:, :, 1::2].repeat_interleave(2, dim=2)[:, :, :te]
    np.testing.assert_allclose(y_back.cpu().numpy(), (y[:, :, :te] * ms).numpy(), rtol=1e-3, atol=2e-4)


def test_fresh_decoder_is_identity_with_zero_logdet():
    """attentions.py:107-110 zero-init + ActNorm zeros + orthogonal W."""
    from glowtts_b200 import models
    torch.manual_seed(0)
    dec = models.FlowSpecDecoder(8, 16, 5, 1, 3, 2, n_split=4, n_sqz=2).cuda()
    y, mask, _, _ = helpers.flow_inputs(helpers.TINY_CFG, 2, 20, 3, False)
    z, ld = dec(y.cuda(), mask.cuda())
    assert ld.abs().max().item() < 1e-3
    # with identity couplings/ActNorm the flow is a product of orthogonal mixes: norms are preserved
    np.testing.assert_allclose(z.detach().pow(2).sum().item(), (y * mask[:, :, 1::2].repeat_interleave(2, 2)).pow(2).sum().item(), rtol=1e-4)


def test_modules_match_reference_fixture(golden_dir):
    from glowtts_b200 import commons, modules
    g = np.load(os.path.join(golden_dir, "modules_small.npz"))
    x = torch.from_numpy(g["x"]).cuda()
    lengths = torch.from_numpy(g["lengths"])
    mask = oflow.sequence_mask(lengths, x.shape[2]).float().unsqueeze(1).cuda()
    an = modules.ActNorm(16).cuda()
    an.logs.data.copy_(torch.from_numpy(g["an_logs"]))
    an.bias.data.copy_(torch.from_numpy(g["an_bias"]))
    z, ld = an(x, mask)
    np.testing.assert_allclose(z.detach().cpu().numpy(), g["an_z"], **TOL)
    np.testing.assert_allclose(ld.detach().cpu().numpy(), g["an_logdet"], **TOL)
    zr, none = an(x, mask, reverse=True)
    assert none is None
    np.testing.assert_allclose(zr.detach().cpu().numpy(), g["an_zrev"], **TOL)
    ddi = modules.ActNorm(16, ddi=True).cuda()
    zd, _ = ddi(x, mask)
    np.testing.assert_allclose(ddi.logs.detach().cpu().numpy(), g["ddi_logs"], **TOL)
    np.testing.assert_allclose(ddi.bias.detach().cpu().numpy(), g["ddi_bias"], **TOL)
    np.testing.assert_allclose(zd.detach().cpu().numpy(), g["ddi_z"], rtol=1e-4, atol=1e-4)
    ic = modules.InvConvNear(16, n_split=4).cuda()
    ic.weight.data.copy_(torch.from_numpy(g["ic_w"]))
    z, ld = ic(x, mask)
    np.testing.assert_allclose(z.detach().cpu().numpy(), g["ic_z"], **TOL)
    np.testing.assert_allclose(ld.detach().cpu().numpy(), g["ic_logdet"], **TOL)
    zr, _ = ic(x, mask, reverse=True)
    np.testing.assert_allclose(zr.detach().cpu().numpy(), g["ic_zrev"], **TOL)
    ic.store_inverse()
    zr2, _ = ic(x, mask, reverse=True)
    np.testing.assert_allclose(zr2.detach().cpu().numpy(), g["ic_zrev"], **TOL)
    out = commons.fused_add_tanh_sigmoid_multiply(torch.from_numpy(g["gate_a"]).cuda(), torch.from_numpy(g["gate_b"]).cuda(),
                                                  torch.IntTensor([16]))
    np.testing.assert_allclose(out.cpu().numpy(), g["gate_out"], **TOL)
    xs, ms = commons.squeeze(x, mask, 2)
    assert np.array_equal(xs.cpu().numpy(), g["sq_x"]) and np.array_equal(ms.cpu().numpy(), g["sq_mask"])
    xu, mu = commons.unsqueeze(xs, ms, 2)
    assert np.array_equal(xu.cpu().numpy(), g["unsq_x"]) and np.array_equal(mu.cpu().numpy(), g["unsq_mask"])
    gp = commons.generate_path(torch.from_numpy(g["gp_dur"]).cuda(), torch.from_numpy(g["gp_mask"]).cuda())
    assert np.array_equal(gp.cpu().numpy(), g["gp_path"])


def test_logp_matches_reference_fixture(golden_dir):
    from glowtts_b200 import commons
    g = np.load(os.path.join(golden_dir, "logp_small.npz"))
    lp = commons.logp(torch.from_numpy(g["x_m"]).cuda(), torch.from_numpy(g["x_logs"]).cuda(), torch.from_numpy(g["z"]).cuda())
    np.testing.assert_allclose(lp.cpu().numpy(), g["logp"], rtol=1e-5, atol=1e-4)


@pytest.mark.parametrize("mean_only", [True, False])
def test_logp_matches_oracle_and_feeds_mas(mean_only):
    from glowtts_b200 import commons, monotonic_align
    from oracle import mas as omas
    gen = torch.Generator().manual_seed(8)
    B, D, TX, TY = 3, 80, 37, 101
    x_m = torch.randn(B, D, TX, generator=gen) * 0.5
    x_logs = torch.zeros(B, D, TX) if mean_only else torch.randn(B, D, TX, generator=gen) * 0.1
    z = torch.randn(B, D, TY, generator=gen)
    want = oflow.logp(x_m, x_logs, z)
    got = commons.logp(x_m.cuda(), None if mean_only else x_logs.cuda(), z.cuda())
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=1e-5, atol=2e-4)
    mask = torch.ones(B, TX, TY)
    path = monotonic_align.maximum_path(got, mask.cuda())
    ref = omas.maximum_path_np(got.cpu().numpy(), mask.numpy())      # bit-exact given the same fp32 value
    assert (path.cpu().numpy() == ref).all()
    ref2 = omas.maximum_path_np(want.numpy(), mask.numpy())
    agree = (path.cpu().numpy().argmax(1) == ref2.argmax(1)).mean()
    assert agree > 0.97                                              # path agreement under 1e-4 logp noise


@pytest.mark.parametrize("kernel", ["umma", "legacy", "ffma"])
@pytest.mark.parametrize("mean_only,B,D,TX,TY", [(True, 3, 80, 180, 860), (False, 2, 80, 37, 101), (True, 2, 80, 129, 258),
                                                (False, 2, 24, 200, 131), (True, 2, 17, 5, 33)])
def test_logp_kernels_match_oracle(kernel, mean_only, B, D, TX, TY):
    """Every logp kernel (tcgen05 kind::tf32 3xTF32 -- the default for mean_only models --, mma.sync 3xTF32, FFMA) against
    the oracle restatement of models.py:1076-1082: tile seams (T_x > 128, T_y > 128), ragged edges (T_y % 4 != 0), k-chunk
    padding (D = 17, 24) and both mean_only settings."""
    from glowtts_b200 import commons
    gen = torch.Generator().manual_seed(18)
    x_m = torch.randn(B, D, TX, generator=gen) * 0.7
    x_logs = torch.zeros(B, D, TX) if mean_only else torch.randn(B, D, TX, generator=gen) * 0.2
    z = torch.randn(B, D, TY, generator=gen)
    want = oflow.logp(x_m.double(), x_logs.double(), z.double())
    env = {"umma": None, "legacy": "GTK_LOGP_LEGACY", "ffma": "GTK_LOGP_FFMA"}[kernel]     # tcgen05 is the default
    if env:
        os.environ[env] = "1"
    try:
        got = commons.logp(x_m.cuda(), None if mean_only else x_logs.cuda(), z.cuda())
    finally:
        if env:
            os.environ.pop(env, None)
    err = float((got.cpu().double() - want).abs().max())
    assert err <= 4e-6 * float(want.abs().max()) + 1e-5, (kernel, err, float(want.abs().max()))


def test_standalone_coupling_block_and_wn_match_oracle():
    from glowtts_b200 import attentions
    cfg = dict(helpers.TINY_CFG, n_blocks=1, gin_channels=4, kernel_size=3, dilation_rate=2, n_layers=3)
    sd = helpers.make_state_dict(cfg, 55)
    cb = attentions.CouplingBlock(16, 16, 3, 2, 3, gin_channels=4).cuda()
    cb.load_state_dict({k[len("flows.2."):]: v for k, v in sd.items() if k.startswith("flows.2.")})
    gen = torch.Generator().manual_seed(3)
    x = torch.randn(2, 16, 19, generator=gen)
    mask = oflow.sequence_mask(torch.tensor([19, 11]), 19).float().unsqueeze(1)
    g = torch.randn(2, 4, 1, generator=gen)
    xr = x.clone().requires_grad_(True)
    sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    z_ref, ld_ref = oflow.coupling(sdr, "flows.2", xr, mask, hidden=16, dilation_rate=2, n_layers=3, g=g)
    (z_ref.pow(2).sum() + ld_ref.sum()).backward()
    xc = x.cuda().requires_grad_(True)
    z, ld = cb(xc, mask.cuda(), g=g.cuda())
    np.testing.assert_allclose(z.detach().cpu().numpy(), z_ref.detach().numpy(), **TOL)
    np.testing.assert_allclose(ld.detach().cpu().numpy(), ld_ref.detach().numpy(), rtol=1e-4, atol=1e-4)
    (z.pow(2).sum() + ld.sum()).backward()
    np.testing.assert_allclose(xc.grad.cpu().numpy(), xr.grad.numpy(), rtol=1e-3, atol=1e-5)
    for name, p in cb.named_parameters():
        if ".wn_pitch." in name or ".wn_energy." in name or name.startswith(("wn_pitch.", "wn_energy.")):
            assert p.grad is None       # identity without a pitch / energy track
            continue
        np.testing.assert_allclose(p.grad.cpu().numpy(), sdr["flows.2." + name].grad.numpy(), rtol=2e-3, atol=2e-5,
                                   err_msg=name)
    with torch.no_grad():
        zr, _ = cb(z.detach(), mask.cuda(), g=g.cuda(), reverse=True)
    np.testing.assert_allclose(zr.cpu().numpy(), x.numpy() * np.concatenate(
        [np.ones((2, 8, 19), np.float32), mask.expand(2, 8, 19).numpy()], 1), rtol=1e-3, atol=1e-4)


def test_mle_loss_and_prior_expansion_match_oracle():
    from glowtts_b200 import commons, ops
    gen = torch.Generator().manual_seed(4)
    B, C, T, TX = 3, 8, 30, 7
    lengths = torch.tensor([30, 22, 16])
    mask = oflow.sequence_mask(lengths, T).float().unsqueeze(1)
    z = torch.randn(B, C, T, generator=gen) * mask
    m = torch.randn(B, C, T, generator=gen) * mask
    logs = torch.randn(B, C, T, generator=gen) * 0.1 * mask
    ld = torch.randn(B, generator=gen)
    ins = [t.clone().requires_grad_(True) for t in (z, m, logs, ld)]
    want = oflow.mle_loss(*ins, mask)
    want.backward()
    gins = [t.cuda().requires_grad_(True) for t in (z, m, logs, ld)]
    got = commons.mle_loss(*gins, mask.cuda())
    got.backward()
    assert abs(got.item() - want.item()) < 1e-5 * max(1, abs(want.item()))
    for a, b in zip(gins, ins):
        np.testing.assert_allclose(a.grad.cpu().numpy(), b.grad.numpy(), rtol=1e-4, atol=1e-7)
    # prior expansion as a gather == attn^T matmul (models.py:1118-1119)
    value, amask, _, _ = helpers.mas_case(B, TX, T, 3)
    v, am = torch.from_numpy(value * amask).cuda(), torch.from_numpy(amask).cuda()
    t_xs, t_ys = ops.mas_lengths(am)
    attn, dur, idx = ops.mas_path(v, t_xs, t_ys)
    x_m = torch.randn(B, C, TX, generator=gen)
    z_m_ref, z_l_ref = oflow.expand_prior(attn.cpu(), x_m, x_m * 2)
    z_m, z_l = commons.expand_prior(x_m.cuda(), (x_m * 2).cuda(), idx)
    assert torch.equal(z_m.cpu(), z_m_ref) and torch.equal(z_l.cpu(), z_l_ref)


# ----------------------------------------------------------------------------------------- dropout (modules.py:120,153)
def _dropout_decoder(cfg, sd, p, precision="fp32"):
    from glowtts_b200 import models
    dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                 cfg["n_blocks"], cfg["n_layers"], p_dropout=p, n_split=cfg["n_split"],
                                 n_sqz=cfg["n_sqz"], gin_channels=cfg["gin_channels"], precision=precision)
    dec.load_state_dict(sd, strict=False)
    return dec.cuda().train()


def test_dropout_mask_statistics_and_determinism():
    from glowtts_b200 import ops
    p = 0.05
    m = ops.dropout_scale(1234, 2, 1, p, 8, 500, 384, "cuda")
    m2 = ops.dropout_scale(1234, 2, 1, p, 8, 500, 384, "cuda")
    other = ops.dropout_scale(1234, 2, 2, p, 8, 500, 384, "cuda")
    assert torch.equal(m, m2)
    assert not torch.equal(m, other)
    vals = torch.unique(m).cpu().numpy()
    thr = round(p * 65536)
    np.testing.assert_allclose(vals, [0.0, 65536.0 / (65536.0 - thr)], rtol=1e-6)
    drop_rate = (m == 0).float().mean().item()
    assert abs(drop_rate - p) < 4 * (p * (1 - p) / m.numel()) ** 0.5 + 1e-5       # 4 sigma
    assert abs(m.mean().item() - 1.0) < 2e-3                                        # unbiased scaling
    # per-channel and per-frame rates are flat (no structure along either axis)
    assert ((m == 0).float().mean(dim=(0, 2)) - p).abs().max().item() < 0.03
    assert ((m == 0).float().mean(dim=(0, 1)) - p).abs().max().item() < 0.03


@pytest.mark.parametrize("use_g", [False, True])
def test_decoder_dropout_matches_oracle_given_the_mask(use_g):
    """Training-mode forward/backward with p_dropout > 0 equals the oracle run with the SAME keep-mask (the mask the
    kernels derive from the per-call seed); eval mode ignores p_dropout."""
    from glowtts_b200 import ops
    p = 0.2
    cfg = dict(helpers.TINY_CFG, n_blocks=2, n_layers=3, hidden_channels=32, gin_channels=5 if use_g else 0)
    sd = helpers.make_state_dict(cfg, 4242)
    dec = _dropout_decoder(cfg, sd, p)
    b, t = 3, 42
    y, mask, _, cond = helpers.flow_inputs(cfg, b, t, 99, use_g)
    torch.manual_seed(777)
    seed = dec.peek_dropout_seed()             # the seed the next training forward uses (device-resident state)
    yc = y.cuda().requires_grad_(True)
    condc = None if cond is None else cond.cuda().requires_grad_(True)
    z, ld = dec(yc, mask.cuda(), g=condc)
    (z.pow(2).sum() + ld.sum()).backward()

    H2 = 2 * cfg["hidden_channels"]
    masks = {(blk, i): ops.dropout_scale(seed, blk, i, p, b, t // 2, H2, "cuda").cpu()
             for blk in range(cfg["n_blocks"]) for i in range(cfg["n_layers"])}
    assert all((m == 0).any() for m in masks.values())
    yr = y.clone().requires_grad_(True)
    condr = None if cond is None else cond.clone().requires_grad_(True)
    sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    z_ref, ld_ref = oflow.decoder(sdr, yr, mask, g=condr, drop=lambda blk, i, v: v * masks[(blk, i)], **_kw(cfg))
    (z_ref.pow(2).sum() + ld_ref.sum()).backward()
    np.testing.assert_allclose(z.detach().cpu().numpy(), z_ref.detach().numpy(), **TOL)
    np.testing.assert_allclose(ld.detach().cpu().numpy(), ld_ref.detach().numpy(), rtol=1e-4, atol=1e-3)
    np.testing.assert_allclose(yc.grad.cpu().numpy(), yr.grad.numpy(), rtol=1e-3, atol=1e-5)
    if use_g:
        np.testing.assert_allclose(condc.grad.cpu().numpy(), condr.grad.numpy(), rtol=2e-3, atol=2e-5)
    n = 0
    for name, prm in dec.named_parameters():
        if name in sdr and sdr[name].grad is not None and prm.grad is not None:
            np.testing.assert_allclose(prm.grad.cpu().numpy().reshape(-1), sdr[name].grad.numpy().reshape(-1), rtol=2e-3,
                                       atol=3e-5, err_msg=name)
            n += 1
    assert n > 20
    # a different call draws a different mask; eval() switches dropout off
    z2, _ = dec(yc.detach(), mask.cuda(), g=None if condc is None else condc.detach())
    assert not torch.allclose(z2, z.detach())
    dec.eval()
    with torch.no_grad():
        z_eval, _ = dec(y.cuda(), mask.cuda(), g=None if cond is None else cond.cuda())
        z_plain, _ = oflow.decoder(sd, y, mask, g=cond, **_kw(cfg))
    np.testing.assert_allclose(z_eval.cpu().numpy(), z_plain.numpy(), **TOL)


@pytest.mark.parametrize("norm_type,clip", [(2, None), (2, 0.05), (1, 0.1), (3, 0.02)])
def test_clip_grad_value_matches_oracle(norm_type, clip):
    """commons.clip_grad_value_ (commons.py:320-336): separate tensors, neighbours in one flat buffer, odd offsets."""
    from glowtts_b200 import commons
    gen = torch.Generator().manual_seed(11)
    flat = torch.randn(10007, generator=gen) * 0.1
    cuts = [0, 3, 1030, 1031, 5000, 10007]                     # views into one buffer, unaligned starts
    loose = [torch.randn(257, 3, generator=gen) * 0.1, torch.randn(1, generator=gen)]
    want_grads = [flat[a:b].clone() for a, b in zip(cuts, cuts[1:])] + [t.clone() for t in loose]
    want = oflow.clip_grad_value_(want_grads, clip, norm_type)
    dflat = flat.cuda()
    params = []
    for a, b in zip(cuts, cuts[1:]):
        p = torch.nn.Parameter(torch.zeros(b - a, device="cuda"))
        p.grad = dflat[a:b]
        params.append(p)
    for t in loose:
        p = torch.nn.Parameter(torch.zeros_like(t, device="cuda"))
        p.grad = t.cuda()
        params.append(p)
    params.append(torch.nn.Parameter(torch.zeros(4, device="cuda")))       # no gradient: skipped
    n0 = _lib_launches()
    got = commons.clip_grad_value_(params[::-1], clip, norm_type)          # any order
    assert _lib_launches() - n0 == 3                                       # one launch for the flat buffer + 2 loose tensors
    assert got.is_cuda and got.dim() == 0
    np.testing.assert_allclose(float(got), want, rtol=2e-5)
    for p, w in zip(params, want_grads):
        assert torch.equal(p.grad.cpu(), w)


def _lib_launches():
    from glowtts_b200 import _lib
    return _lib.launch_count()
