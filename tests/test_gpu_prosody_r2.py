"""Round 2: the live fork's 3-WaveNet coupling block (wn -> wn_energy -> wn_pitch,
attentions.py:155-160, modules.WNP modules.py:272-362) INSIDE the fused bf16 tcgen05 decoder call, against fixtures made
by the reference's own modules at tensor-core-capable widths (tests/golden/make_golden.py --prosody-h64-only), with the
same bounds as test_gpu_parity_r2.py; the fp32 precision runs the same fixtures through the module-level ops."""
import os

import numpy as np
import pytest
import torch

import helpers
from test_gpu_parity_r2 import _check_against, _rel_l2

pytestmark = pytest.mark.gpu

H64_PE = dict(in_channels=16, hidden_channels=64, kernel_size=3, dilation_rate=2, n_blocks=2, n_layers=3, n_split=4,
              n_sqz=2, gin_channels=8)
H64_P = dict(in_channels=16, hidden_channels=64, kernel_size=5, dilation_rate=1, n_blocks=2, n_layers=3, n_split=4,
             n_sqz=2, gin_channels=0)
CASES = {"flow_h64_pe": H64_PE, "flow_h64_p": H64_P}


def _decoder(cfg, sd, precision, p_dropout=0.0):
    from glowtts_b200 import models
    dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                 cfg["n_blocks"], cfg["n_layers"], p_dropout=p_dropout, n_split=cfg["n_split"],
                                 n_sqz=cfg["n_sqz"], gin_channels=cfg["gin_channels"], precision=precision)
    dec.load_state_dict(sd, strict=True)
    return dec.cuda()


def _inputs(cfg, g):
    y, mask, _, cond = helpers.flow_inputs(cfg, int(g["b"]), int(g["t"]), int(g["in_seed"]), bool(g["use_g"]))
    pitch, energy = helpers.prosody_inputs(int(g["b"]), int(g["t"]), int(g["in_seed"]), bool(g["use_pitch"]),
                                           bool(g["use_energy"]))
    cu = lambda t: None if t is None else t.cuda()
    return cu(y), cu(mask), cu(cond), cu(pitch), cu(energy)


@pytest.mark.parametrize("name,precision", [(n, p) for n in CASES for p in ("fp32", "bf16")])
def test_prosody_decoder_vs_reference_fixture(golden_dir, name, precision):
    from glowtts_b200 import _lib, commons
    cfg = CASES[name]
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    dec = _decoder(cfg, helpers.make_state_dict(cfg, int(g["sd_seed"])), precision)
    y, mask, cond, pitch, energy = _inputs(cfg, g)
    y.requires_grad_(True)
    pitch.requires_grad_(True)
    if energy is not None:
        energy.requires_grad_(True)
    n0 = _lib.launch_count()
    z, ld = dec(y, mask, g=cond, pitch=pitch, energy=energy)
    n_fwd = _lib.launch_count() - n0
    loss = commons.mle_loss(z, None, None, ld, mask)
    loss.backward()
    with torch.no_grad():
        y_rev, none = dec(z.detach(), mask, g=cond, pitch=pitch, energy=energy, reverse=True)
    assert none is None
    grads = {k: p.grad for k, p in dec.named_parameters() if p.grad is not None}
    want = dict(z=g["z"], logdet=g["logdet"], loss=float(g["loss"]), dy=g["dy"], y_rev=g["y_rev"],
                grads={k[5:]: g[k] for k in g.files if k.startswith("grad:")})
    # the chained WaveNets' own tensors are pinned too, cond_layer1 included
    assert any(".wn_pitch.cond_layer1.weight_g" in k for k in want["grads"])
    assert (".wn_energy." in " ".join(want["grads"])) == bool(g["use_energy"])
    for k in want["grads"]:
        assert k in grads, f"no gradient for {k}"
    _check_against(name, precision, z, ld, loss.item(), y.grad, grads, want, y_rev, y)
    tol = 1e-3 if precision == "fp32" else 3e-2
    e = _rel_l2(pitch.grad.cpu().reshape(g["dpitch"].shape), torch.as_tensor(g["dpitch"]))
    assert e < tol, ("dpitch", e)
    if energy is not None:
        e = _rel_l2(energy.grad.cpu().reshape(g["denergy"].shape), torch.as_tensor(g["denergy"]))
        assert e < tol, ("denergy", e)
    if precision == "bf16":
        # ONE fused call: a flow-by-flow composition of the three WaveNets takes several hundred launches at this size
        L, NB, n_wn = cfg["n_layers"], cfg["n_blocks"], 2 + int(bool(g["use_energy"]))
        assert n_fwd <= NB * (n_wn * L + 6) + 16, n_fwd
    # store_inverse(): folded weights (attentions.py:188-194 folds all three WaveNets)
    dec.store_inverse()
    with torch.no_grad():
        y_rev2, _ = dec(z.detach(), mask, g=cond, pitch=pitch, energy=energy, reverse=True)
    w = torch.as_tensor(g["y_rev_stored"])
    err = float((y_rev2.cpu() - w).abs().max() / w.abs().max())
    assert err < (1e-4 if precision == "fp32" else 5e-2), err


def test_prosody_fused_matches_module_path_ragged_batch():
    """Fused bf16 chain against this library's own fp32 module-level composition on a ragged batch that spans several
    128-frame tiles per utterance (packing gaps, tile-pair early exit), tracks given as [B,T] and longer than T_y."""
    from glowtts_b200 import commons
    cfg = H64_PE
    sd = helpers.make_state_dict(cfg, 61)
    B, T = 5, 700
    gen = torch.Generator().manual_seed(62)
    lengths = torch.tensor([700, 514, 2, 388, 130])
    mask = (torch.arange(T)[None, :] < lengths[:, None]).float().unsqueeze(1).cuda()
    y = (torch.randn(B, cfg["in_channels"], T, generator=gen).cuda() * mask)
    cond = torch.randn(B, cfg["gin_channels"], 1, generator=gen).cuda()
    pitch = torch.randn(B, T + 6, generator=gen).cuda()
    energy = torch.randn(B, 1, T, generator=gen).cuda()
    out = {}
    for prec in ("fp32", "bf16"):
        dec = _decoder(cfg, sd, prec)
        yy = y.clone().requires_grad_(True)
        pp, ee = pitch.clone().requires_grad_(True), energy.clone().requires_grad_(True)
        z, ld = dec(yy, mask, g=cond, pitch=pp[:, :T] if prec == "fp32" else pp, energy=ee)
        loss = commons.mle_loss(z, None, None, ld, mask)
        loss.backward()
        out[prec] = (z.detach(), ld.detach(), loss.item(), yy.grad, pp.grad, ee.grad,
                     {k: p.grad for k, p in dec.named_parameters() if p.grad is not None})
    a, b = out["bf16"], out["fp32"]
    assert torch.isfinite(a[0]).all()
    assert float((a[0] - b[0]).abs().max() / b[0].abs().max()) < 2e-2
    assert float(((a[1] - b[1]).abs() / (b[1].abs() * 2e-2 + 0.5)).max()) < 1.0
    assert abs(a[2] - b[2]) < 5e-3 * max(1.0, abs(b[2]))
    assert _rel_l2(a[3], b[3]) < 3e-2
    assert a[4].shape == pitch.shape and float(a[4][:, T:].abs().max()) == 0.0      # frames beyond T_y: no gradient
    assert _rel_l2(a[4], b[4]) < 3e-2 and _rel_l2(a[5], b[5]) < 3e-2
    # padded frames of every utterance get exactly zero track gradient
    assert float((a[4][:, :T] * (1 - mask[:, 0])).abs().max()) == 0.0
    # cond_layer1.weight_v: weight norm over ONE element, w = g sign(v), so d/dv is analytically 0 (rounding noise in fp32)
    skip = lambda k: k.endswith("cond_layer1.weight_v")
    for k in b[6]:
        if skip(k):
            assert float(a[6][k].abs().max()) <= 1e-6 * float(b[6][k.replace("weight_v", "weight_g")].abs().max()), k
    worst = max((_rel_l2(a[6][k].float(), b[6][k]), k) for k in b[6] if b[6][k].norm() > 1e-7 and not skip(k))
    print("[prosody fused vs module path] worst grad relL2", worst)
    assert worst[0] < 3e-2, worst


@pytest.mark.parametrize("use_energy", [True, False])
def test_prosody_dropout_matches_oracle_given_the_mask(use_energy):
    """p_dropout > 0 through the chained WaveNets (modules.py:153, 331): forward and every gradient -- cond_layer1 and
    the tracks take the gate-input gradient BEFORE the dropout scaling -- equal the oracle run with the keep-masks the
    kernels derive from the per-call seed (layer index = WaveNet id x n_layers + layer)."""
    from glowtts_b200 import ops
    from oracle import flow as oflow
    p = 0.2
    cfg = dict(H64_P, n_layers=2)
    L, H2 = cfg["n_layers"], 2 * cfg["hidden_channels"]
    sd = helpers.make_state_dict(cfg, 81)
    dec = _decoder(cfg, sd, "bf16", p_dropout=p).train()
    b, t = 3, 300
    y, mask, _, _ = helpers.flow_inputs(cfg, b, t, 82, False)
    pitch, energy = helpers.prosody_inputs(b, t, 83, True, use_energy)
    torch.manual_seed(777)
    seed = dec.peek_dropout_seed()
    yc, pc = y.cuda().requires_grad_(True), pitch.cuda().requires_grad_(True)
    ec = None if energy is None else energy.cuda().requires_grad_(True)
    z, ld = dec(yc, mask.cuda(), pitch=pc, energy=ec)
    (z.pow(2).sum() + ld.sum()).backward()
    masks = {(blk, i): ops.dropout_scale(seed, blk, i, p, b, t // 2, H2, "cuda").cpu()
             for blk in range(cfg["n_blocks"]) for i in range(3 * L)}
    yr, pr = y.clone().requires_grad_(True), pitch.clone().requires_grad_(True)
    er = None if energy is None else energy.clone().requires_grad_(True)
    sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    z_ref, ld_ref = oflow.decoder(sdr, yr, mask, pitch=pr, energy=er, drop=lambda blk, i, v: v * masks[(blk, i)],
                                  hidden=cfg["hidden_channels"], dilation_rate=cfg["dilation_rate"], n_blocks=cfg["n_blocks"],
                                  n_layers=L, n_split=cfg["n_split"], n_sqz=cfg["n_sqz"])
    (z_ref.pow(2).sum() + ld_ref.sum()).backward()
    assert float((z.detach().cpu() - z_ref.detach()).abs().max() / z_ref.detach().abs().max()) < 2e-2
    assert _rel_l2(yc.grad.cpu(), yr.grad) < 3e-2
    assert _rel_l2(pc.grad.cpu(), pr.grad) < 3e-2
    if use_energy:
        assert _rel_l2(ec.grad.cpu(), er.grad) < 3e-2
    n, worst = 0, (0.0, "")
    for name, prm in dec.named_parameters():
        if (name in sdr and sdr[name].grad is not None and prm.grad is not None and sdr[name].grad.norm() > 1e-7
                and not name.endswith("cond_layer1.weight_v")):
            worst = max(worst, (_rel_l2(prm.grad.cpu().reshape(-1), sdr[name].grad.reshape(-1)), name))
            n += 1
    print("[prosody dropout vs oracle] tensors", n, "worst grad relL2", worst)
    assert n > 30 and worst[0] < 3e-2, worst
    assert any("wn_pitch.cond_layer1" in k for k, q in dec.named_parameters() if q.grad is not None)


def test_prosody_dropout_training_step_graphed():
    """p_dropout > 0 (the gate-input gradient is kept un-scaled for cond_layer1 / the tracks), flat parameters with the
    chained WaveNets inside, the whole step as a CUDA graph: finite, reproducible from the seed, and it trains."""
    from glowtts_b200 import hot_path
    cfg = dict(H64_PE, gin_channels=0)
    B, T, Tx = 4, 512, 40
    gen = torch.Generator().manual_seed(71)
    y = torch.randn(B, cfg["in_channels"], T, generator=gen)
    y_len = torch.tensor([512, 300, 450, 128])
    x_m = torch.randn(B, cfg["in_channels"], Tx, generator=gen)
    x_len = torch.tensor([40, 25, 33, 12])
    pitch = torch.rand(B, 1, T, generator=gen) * 200 + 80
    pitch[:, :, ::7] = 0.0                                         # unvoiced frames stay exactly zero (models.py:1060-1063)
    energy = torch.rand(B, 1, T, generator=gen) + 0.1
    batch = [t.cuda() for t in (y, y_len, x_m, x_len, pitch, energy)]
    losses = []
    for rep in range(2):
        torch.manual_seed(5)
        dec = _decoder(cfg, helpers.make_state_dict(cfg, 72), "bf16", p_dropout=0.05).train()
        params = dec.flat_parameters(prosody=True)
        assert len(params) == 1                                    # everything the fused call trains is in the flat buffer
        opt = torch.optim.AdamW(params, lr=2e-4, capturable=True)
        step = hot_path.GraphedTrainStep(dec, opt, batch)
        ls = [float(step.step().detach()) for _ in range(6)]
        losses.append(ls)
        g = dec.flows[2].wn_pitch.cond_layer1.weight_g.grad
        assert g is not None and torch.isfinite(g).all() and float(g.abs().max()) > 0
        assert dec.flows[2].wn_energy.in_layers[0].weight_v.grad is not None
    assert all(np.isfinite(l) for l in losses[0])
    # same seeds -> same dropout masks -> same trajectory (up to the summation order of the atomic reductions)
    assert np.allclose(losses[0], losses[1], rtol=1e-5), losses
    assert losses[0][-1] < losses[0][0], losses[0]


def test_all_taps_weight_gradient_tiles_match_per_tap(tmp_path):
    """GTK_WGRAD_ALL_TAPS=1 (wgrad_group.cu: every tap of a k > 1 conv inside one tile, row-shifted MN-major descriptors on
    one staged X tile) gives the same parameter gradients as the default per-tap tiles.  The switch is read once per
    process, so each mode runs in its own interpreter."""
    import subprocess
    import sys
    script = r'''
import sys, numpy as np, torch
sys.path.insert(0, "tests")
import helpers
from glowtts_b200 import models, commons
cfg = dict(in_channels=16, hidden_channels=192, kernel_size=5, dilation_rate=1, n_blocks=2, n_layers=2, n_split=4, n_sqz=2, gin_channels=0)
dec = models.FlowSpecDecoder(16, 192, 5, 1, 2, 2, p_dropout=0.0, n_split=4, n_sqz=2, precision="bf16")
dec.load_state_dict(helpers.make_state_dict(cfg, 91), strict=True)
dec = dec.cuda()
gen = torch.Generator().manual_seed(92)
B, T = 3, 900
lengths = torch.tensor([900, 640, 258])
mask = (torch.arange(T)[None, :] < lengths[:, None]).float().unsqueeze(1).cuda()
y = torch.randn(B, 16, T, generator=gen).cuda() * mask
z, ld = dec(y, mask)
commons.mle_loss(z, None, None, ld, mask).backward()
np.savez(sys.argv[1], **{k: p.grad.cpu().numpy() for k, p in dec.named_parameters() if p.grad is not None})
'''
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    # per-tap tiles | tap-pair tiles (the default: two taps share one staged X tile) | all-taps tiles
    for i, mode in enumerate((dict(GTK_WGRAD_TAP_PAIRS="0"), dict(), dict(GTK_WGRAD_ALL_TAPS="1"))):
        out = str(tmp_path / f"g{i}.npz")
        env = dict(os.environ, **mode)
        subprocess.run([sys.executable, "-c", script, out], check=True, cwd=root, env=env, timeout=600)
        outs.append(np.load(out))
    a = outs[0]
    for b in outs[1:]:
        assert set(a.files) == set(b.files) and len(a.files) > 20
        for k in a.files:
            ref = np.abs(a[k]).max()
            assert np.abs(a[k] - b[k]).max() <= 2e-3 * ref + 1e-7, (k, float(np.abs(a[k] - b[k]).max()), float(ref))
