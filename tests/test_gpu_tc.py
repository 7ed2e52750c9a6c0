"""GPU checks of the bf16 tcgen05 path against torch fp32 references of the same op evaluated on the
SAME bf16-rounded operands (so the only difference is fp32 accumulation order): rtol 2e-3."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


@pytest.mark.parametrize("B,Cin,M,T,K,dil", [(2, 192, 384, 430, 5, 1), (3, 80, 192, 77, 1, 1), (1, 192, 160, 128, 1, 1),
                                              (2, 192, 192, 300, 3, 2), (2, 384, 192, 129, 5, 1), (1, 64, 16, 5, 5, 3)])
def test_tc_conv1d_matches_torch(B, Cin, M, T, K, dil):
    from glowtts_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(B * 1000 + T)
    x = torch.randn(B, Cin, T, device="cuda", generator=gen)
    w = torch.randn(M, Cin, K, device="cuda", generator=gen) / (Cin * K) ** 0.5
    bias = torch.randn(M, device="cuda", generator=gen)
    lengths = torch.randint(T // 2, T + 1, (B,), device="cuda", generator=gen)
    mask = (torch.arange(T, device="cuda")[None, :] < lengths[:, None]).float()
    frames = ops.tc_frames(x)
    assert frames.shape == (B, T, (Cin + 63) // 64 * 64)
    np.testing.assert_array_equal(frames[:, :, :Cin].float().cpu().numpy(), _bf(x).permute(0, 2, 1).cpu().numpy())
    assert frames[:, :, Cin:].abs().sum().item() == 0
    wp = ops.tc_pack_weight(w)
    out32, out16 = ops.tc_conv1d(frames, wp, bias, dil, mask, want_f32=True, want_bf16=True)
    ref = F.conv1d(_bf(x), _bf(w), bias, padding=(K - 1) * dil // 2, dilation=dil) * mask[:, None, :]
    ref = ref.permute(0, 2, 1)
    torch.cuda.synchronize()
    np.testing.assert_allclose(out32.cpu().numpy(), ref.cpu().numpy(), rtol=2e-3, atol=2e-3)
    np.testing.assert_allclose(out16[:, :, :M].float().cpu().numpy(), ref.cpu().numpy(), rtol=1e-2, atol=1e-2)


def test_tc_conv1d_data_gradient_form():
    """conv^T: the packed transposed+flipped weights give the gradient w.r.t. the conv input."""
    from glowtts_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(7)
    B, Cin, M, T, K, dil = 2, 192, 384, 200, 5, 1
    x = torch.randn(B, Cin, T, device="cuda", generator=gen, requires_grad=True)
    w = torch.randn(M, Cin, K, device="cuda", generator=gen) / (Cin * K) ** 0.5
    dout = torch.randn(B, M, T, device="cuda", generator=gen)
    F.conv1d(x, _bf(w), None, padding=(K - 1) * dil // 2, dilation=dil).backward(_bf(dout))
    wpt = ops.tc_pack_weight(w, transpose=True)
    dx, _ = ops.tc_conv1d(ops.tc_frames(dout), wpt, None, dil)
    np.testing.assert_allclose(dx.cpu().numpy(), x.grad.permute(0, 2, 1).cpu().numpy(), rtol=2e-3, atol=2e-3)


@pytest.mark.parametrize("last,use_cond", [(False, True), (True, False), (False, False)])
def test_fused_wn_layer_matches_torch(last, use_cond):
    """One WaveNet layer (modules.py:151-170) in the fused tcgen05 kernel vs torch on bf16-rounded operands."""
    import ctypes
    from glowtts_b200 import _lib, ops
    gen = torch.Generator(device="cuda").manual_seed(11 + int(last))
    B, H, T, K, dil = 3, 192, 301, 5, 1
    RS = H if last else 2 * H
    h = torch.randn(B, H, T, device="cuda", generator=gen)
    lengths = torch.tensor([T, T - 40, T // 2], device="cuda")
    mask = (torch.arange(T, device="cuda")[None, :] < lengths[:, None]).float()
    h = h * mask[:, None]
    win = torch.randn(2 * H, H, K, device="cuda", generator=gen) / (H * K) ** 0.5
    wrs = torch.randn(RS, H, 1, device="cuda", generator=gen) / H ** 0.5
    b_in = torch.randn(2 * H, device="cuda", generator=gen) * 0.1
    b_rs = torch.randn(RS, device="cuda", generator=gen) * 0.1
    cond = torch.randn(B, 2 * H, device="cuda", generator=gen) * 0.3 if use_cond else None
    skip_prev = torch.randn(B, T, H, device="cuda", generator=gen)
    # reference
    xin = F.conv1d(_bf(h), _bf(win), b_in, padding=(K - 1) * dil // 2, dilation=dil)
    if cond is not None:
        xin = xin + cond[:, :, None]
    tt, ss = torch.tanh(xin[:, :H]), torch.sigmoid(xin[:, H:])
    acts = tt * ss
    rs = F.conv1d(_bf(acts), _bf(wrs), b_rs)
    # kernel
    hf = h.permute(0, 2, 1).contiguous()
    hb = hf.to(torch.bfloat16)
    win_p, wrs_p = ops.tc_pack_weight(win), ops.tc_pack_weight(wrs)
    h_out32 = torch.zeros(B, T, H, device="cuda")
    h_out16 = torch.zeros(B, T, H, device="cuda", dtype=torch.bfloat16)
    skip32 = skip_prev.clone()
    skip16 = torch.zeros(B, T, H, device="cuda", dtype=torch.bfloat16)
    ts = torch.zeros(B, T, 2 * H, device="cuda", dtype=torch.bfloat16)
    ac = torch.zeros(B, T, H, device="cuda", dtype=torch.bfloat16)
    P = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
    _lib.call("gtk_tc_wn_layer", P(hb), P(hf), P(win_p), P(wrs_p), P(b_in), P(b_rs), P(cond), 2 * H, P(mask), B, T, H, RS, K, dil,
              0, int(last), ctypes.c_float(0.0), 0, P(h_out32), P(h_out16), P(skip32), P(skip16), P(ts), P(ac),
              ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    tol = dict(rtol=1e-2, atol=1e-2)
    # Masked frames: the reference's values there never reach an output (x * mask, output * mask: modules.py:166,171);
    # the kernel writes exact zeros for them, so the comparison is over the valid frames (and zeros are checked below).
    mk3 = mask[:, :, None]
    np.testing.assert_allclose(ts[:, :, :H].float().cpu().numpy(), (tt.permute(0, 2, 1) * mk3).cpu().numpy(), **tol)
    np.testing.assert_allclose(ts[:, :, H:].float().cpu().numpy(), (ss.permute(0, 2, 1) * mk3).cpu().numpy(), **tol)
    np.testing.assert_allclose(ac.float().cpu().numpy(), (acts.permute(0, 2, 1) * mk3).cpu().numpy(), **tol)
    if last:
        want = (skip_prev + rs.permute(0, 2, 1)) * mask[:, :, None]
        np.testing.assert_allclose(skip16.float().cpu().numpy(), want.cpu().numpy(), rtol=2e-2, atol=2e-2)
    else:
        want_h = (h + rs[:, :H]).permute(0, 2, 1) * mask[:, :, None]
        np.testing.assert_allclose(h_out32.cpu().numpy(), want_h.cpu().numpy(), rtol=5e-3, atol=5e-3)
        np.testing.assert_allclose(h_out16.float().cpu().numpy(), want_h.cpu().numpy(), rtol=2e-2, atol=2e-2)
        want_s = skip_prev + rs[:, H:].permute(0, 2, 1)
        np.testing.assert_allclose((skip32 * mk3).cpu().numpy(), (want_s * mk3).cpu().numpy(), rtol=5e-3, atol=5e-3)


def _decoders(cfg, seed, end_std=0.05):
    import helpers
    from glowtts_b200 import models
    sd = helpers.make_state_dict(cfg, seed, end_std=end_std)
    out = []
    for prec in ("fp32", "bf16"):
        dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                     cfg["n_blocks"], cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"],
                                     gin_channels=cfg["gin_channels"], precision=prec)
        dec.load_state_dict(sd, strict=False)
        out.append(dec.cuda())
    return out


TC_CFGS = {
    "h64": dict(in_channels=16, hidden_channels=64, kernel_size=5, dilation_rate=1, n_blocks=3, n_layers=3, n_split=4,
                n_sqz=2, gin_channels=0),
    "h128_g_dil2": dict(in_channels=16, hidden_channels=128, kernel_size=3, dilation_rate=2, n_blocks=2, n_layers=2,
                        n_split=4, n_sqz=2, gin_channels=8),
    "base": dict(in_channels=80, hidden_channels=192, kernel_size=5, dilation_rate=1, n_blocks=12, n_layers=4, n_split=4,
                 n_sqz=2, gin_channels=0),
}


@pytest.mark.parametrize("name", list(TC_CFGS))
def test_bf16_decoder_forward_and_reverse_track_fp32(name):
    """Stated bound for the bf16 WN path (BASELINE.json): max |dz| <= 2e-2 * max|z|, |d logdet| <= 2e-2 * |logdet| + 0.5,
    |d mle_loss| <= 5e-3, and reverse(forward(y)) == y within 5e-2 * max|y|."""
    import helpers
    from glowtts_b200 import commons
    cfg = TC_CFGS[name]
    f32, b16 = _decoders(cfg, 77, end_std=0.01 if name == "base" else 0.05)
    B, T = (4, 300) if name != "base" else (3, 214)
    y, mask, _, g = helpers.flow_inputs(cfg, B, T, 5, cfg["gin_channels"] > 0)
    y, mask = y.cuda(), mask.cuda()
    g = g.cuda() if g is not None else None
    with torch.no_grad():
        z0, ld0 = f32(y, mask, g=g)
        z1, ld1 = b16(y, mask, g=g)
        l0 = commons.mle_loss(z0, None, None, ld0, mask).item()
        l1 = commons.mle_loss(z1, None, None, ld1, mask).item()
        yb, _ = b16(z1, mask, g=g, reverse=True)
    zerr = (z1 - z0).abs().max().item() / z0.abs().max().item()
    lderr = ((ld1 - ld0).abs() / (ld0.abs() * 2e-2 + 0.5)).max().item()
    rerr = (yb - y).abs().max().item() / y.abs().max().item()
    print(f"{name}: rel|dz|={zerr:.2e} logdet={lderr:.2e} dloss={abs(l1 - l0):.2e} roundtrip={rerr:.2e}")
    assert zerr < 2e-2 and lderr < 1.0 and abs(l1 - l0) < 5e-3 and rerr < 5e-2


@pytest.mark.parametrize("B,T,M,Cin,K,dil", [(4, 300, 384, 192, 5, 1), (2, 100, 192, 80, 1, 1), (3, 77, 160, 192, 1, 1),
                                              (2, 215, 128, 64, 3, 2)])
def test_tc_wgrad_and_colsum_match_torch(B, T, M, Cin, K, dil):
    import ctypes
    from glowtts_b200 import _lib, ops
    gen = torch.Generator(device="cuda").manual_seed(T)
    x = torch.randn(B, Cin, T, device="cuda", generator=gen, requires_grad=False)
    dy = torch.randn(B, M, T, device="cuda", generator=gen)
    w = torch.zeros(M, Cin, K, device="cuda", requires_grad=True)
    F.conv1d(_bf(x), w, None, padding=(K - 1) * dil // 2, dilation=dil).backward(_bf(dy))
    xf, dyf = ops.tc_frames(x), ops.tc_frames(dy)
    dw = torch.zeros(M, Cin, K, device="cuda")
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.call("gtk_tc_wgrad", P(dyf), dyf.shape[2], P(xf), xf.shape[2], P(dw), B, T, M, Cin, K, dil, st)
    torch.cuda.synchronize()
    scale = w.grad.abs().max().item()
    np.testing.assert_allclose(dw.cpu().numpy(), w.grad.cpu().numpy(), rtol=2e-3, atol=2e-3 * scale)
    out = torch.zeros(B, M, device="cuda")
    _lib.call("gtk_tc_colsum", P(dyf), dyf.shape[2], M, B, T, P(out), M, st)
    np.testing.assert_allclose(out.cpu().numpy(), _bf(dy).sum(2).cpu().numpy(), rtol=1e-3, atol=1e-3)
    tot = torch.zeros(M, device="cuda")
    _lib.call("gtk_tc_colsum", P(dyf), dyf.shape[2], M, 1, B * T, P(tot), 0, st)
    np.testing.assert_allclose(tot.cpu().numpy(), _bf(dy).sum((0, 2)).cpu().numpy(), rtol=1e-3, atol=2e-3)


@pytest.mark.parametrize("name", list(TC_CFGS))
def test_bf16_decoder_backward_tracks_fp32(name):
    """Gradients of the bf16 path vs the fp32 path (same loss): per-tensor relative L2 error <= 3e-2
    (bf16 operand rounding through up to 48 layers), dy relative L2 error <= 2e-2."""
    import helpers
    from glowtts_b200 import commons
    cfg = TC_CFGS[name]
    f32, b16 = _decoders(cfg, 78, end_std=0.01 if name == "base" else 0.05)
    B, T = (4, 300) if name != "base" else (3, 214)
    y, mask, _, g = helpers.flow_inputs(cfg, B, T, 6, cfg["gin_channels"] > 0)
    mask = mask.cuda()
    g = g.cuda() if g is not None else None
    grads = []
    for dec in (f32, b16):
        yy = y.cuda().requires_grad_(True)
        gg = g.clone().requires_grad_(True) if g is not None else None
        z, ld = dec(yy, mask, g=gg)
        loss = commons.mle_loss(z, None, None, ld, mask)
        loss.backward()
        d = {k: p.grad.clone() for k, p in dec.named_parameters() if p.grad is not None}
        d["__dy"] = yy.grad.clone()
        if gg is not None:
            d["__dg"] = gg.grad.clone()
        grads.append(d)
    worst = ("", 0.0)
    for k, a in grads[0].items():
        b = grads[1][k]
        assert torch.isfinite(b).all(), k
        err = ((a - b).norm() / (a.norm() + 1e-12)).item()
        if a.norm().item() > 1e-7 and err > worst[1]:
            worst = (k, err)
    print(f"{name}: worst relative L2 grad error {worst[1]:.3e} at {worst[0]}; dy {((grads[0]['__dy'] - grads[1]['__dy']).norm() / grads[0]['__dy'].norm()).item():.3e}")
    assert worst[1] < 3e-2


@pytest.mark.parametrize("use_g", [False, True])
def test_bf16_decoder_dropout_matches_fp32_path_with_the_same_seed(use_g):
    """p_dropout > 0: the tcgen05 path and the fp32 path draw the same keep-mask from the same seed, so forward and
    gradients agree within the bf16 bound (the dropped elements are exactly the same ones)."""
    import helpers
    from glowtts_b200 import models
    cfg = dict(helpers.TINY_CFG, in_channels=16, n_blocks=2, n_layers=2, hidden_channels=64, gin_channels=4 if use_g else 0)
    sd = helpers.make_state_dict(cfg, 31)
    decs = {}
    for prec in ("fp32", "bf16"):
        d = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                   cfg["n_blocks"], cfg["n_layers"], p_dropout=0.25, n_split=cfg["n_split"],
                                   n_sqz=cfg["n_sqz"], gin_channels=cfg["gin_channels"], precision=prec)
        d.load_state_dict(sd, strict=False)
        decs[prec] = d.cuda().train()
    y, mask, _, cond = helpers.flow_inputs(cfg, 3, 300, 5, use_g)
    res = {}
    for prec, d in decs.items():
        torch.manual_seed(2024)                # both decoders start their dropout state from the same draw
        yc = y.cuda().requires_grad_(True)
        cc = None if cond is None else cond.cuda().requires_grad_(True)
        z, ld = d(yc, mask.cuda(), g=cc)
        (z.pow(2).mean() + ld.mean() * 1e-3).backward()
        grads = {n: p.grad.detach().clone() for n, p in d.named_parameters() if p.grad is not None}
        res[prec] = (z.detach(), ld.detach(), yc.grad.clone(), None if cc is None else cc.grad.clone(), grads)
    zf, ldf, dyf, dcf, gf = res["fp32"]
    zb, ldb, dyb, dcb, gb = res["bf16"]

    def rel(a, b):
        return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()
    assert rel(zb, zf) < 2e-2, rel(zb, zf)
    assert rel(ldb, ldf) < 2e-2
    assert rel(dyb, dyf) < 3e-2, rel(dyb, dyf)
    if use_g:
        assert rel(dcb, dcf) < 5e-2, rel(dcb, dcf)
    worst = max(rel(gb[n], gf[n]) for n in gf if gf[n].norm() > 1e-6)
    assert worst < 6e-2, worst
    # the un-dropped forward differs by far more than the bf16 bound: the masks really were applied and shared
    decs["bf16"].eval()
    with torch.no_grad():
        z_eval, _ = decs["bf16"](y.cuda(), mask.cuda(), g=None if cond is None else cond.cuda())
    assert rel(z_eval, zf) > 3 * rel(zb, zf)


def test_flat_parameters_and_graphed_step_match_eager_training():
    """Two optimizer steps: (a) eager, per-parameter AdamW; (b) flat-parameter AdamW; (c) two replays of the step captured
    as a CUDA graph (hot_path.GraphedTrainStep: its warm-up passes are rolled back, capturing itself executes nothing).  Same seeds -> same dropout masks -> the losses and the updated weights
    agree (the graph replays exactly the eager kernels; flat vs per-parameter AdamW is the same elementwise math)."""
    import helpers
    from glowtts_b200 import models
    from glowtts_b200.hot_path import GraphedTrainStep, train_hot_path
    cfg = dict(helpers.TINY_CFG, in_channels=16, n_blocks=2, n_layers=2, hidden_channels=64)
    sd = helpers.make_state_dict(cfg, 77)
    b, t_y, t_x = 3, 96, 20
    gen = torch.Generator().manual_seed(5)
    y = torch.randn(b, cfg["in_channels"], t_y, generator=gen)
    y_len = torch.tensor([96, 64, 80])
    x_m = torch.randn(b, cfg["in_channels"], t_x, generator=gen)
    x_len = torch.tensor([20, 11, 15])
    batch = [v.cuda() for v in (y, y_len, x_m, x_len)]

    def make(mode):
        torch.manual_seed(99)
        dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                     cfg["n_blocks"], cfg["n_layers"], p_dropout=0.1, n_split=cfg["n_split"],
                                     n_sqz=cfg["n_sqz"], precision="bf16")
        dec.load_state_dict(sd, strict=False)
        dec = dec.cuda().train()
        prm = dec.parameters() if mode == "eager" else dec.flat_parameters()
        opt = torch.optim.AdamW(prm, lr=1e-3, fused=True, capturable=(mode == "graph"))
        return dec, opt

    losses, weights = {}, {}
    for mode in ("eager", "flat", "graph"):
        dec, opt = make(mode)
        ls = []
        if mode == "graph":
            stepper = GraphedTrainStep(dec, opt, batch, warmup=1)    # warm-up is snapshot/restored: it trains nothing
            ls.append(stepper.step(batch).item())                    # step 1 = first replay
            ls.append(stepper.step(batch).item())
        else:
            for _ in range(2):
                opt.zero_grad(set_to_none=True)
                out = train_hot_path(dec, *batch[:3], None, batch[3])
                out.loss.backward()
                opt.step()
                ls.append(out.loss.item())
        losses[mode] = ls
        weights[mode] = {k: v.detach().clone() for k, v in dec.state_dict().items()}
    assert losses["eager"] == pytest.approx(losses["flat"], rel=1e-5, abs=1e-6)
    assert losses["graph"] == pytest.approx(losses["eager"], rel=1e-5, abs=1e-6)
    for k in weights["eager"]:
        assert torch.allclose(weights["eager"][k], weights["flat"][k], rtol=1e-5, atol=1e-7), k
        assert torch.allclose(weights["eager"][k], weights["graph"][k], rtol=1e-5, atol=1e-7), k


def test_graphed_reverse_matches_eager_reverse():
    """hot_path.GraphedReverse: the inference call z -> mel replayed as one CUDA graph returns what the eager call does."""
    import helpers
    from glowtts_b200 import models
    from glowtts_b200.hot_path import GraphedReverse
    cfg = dict(helpers.TINY_CFG, in_channels=16, n_blocks=3, n_layers=2, hidden_channels=64, gin_channels=4)
    sd = helpers.make_state_dict(cfg, 61)
    for precision in ("fp32", "bf16"):
        dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                     cfg["n_blocks"], cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"],
                                     gin_channels=4, precision=precision)
        dec.load_state_dict(sd, strict=True)
        dec = dec.cuda().eval()
        dec.store_inverse()
        y, mask, _, g = helpers.flow_inputs(cfg, 3, 40, 9, True)
        y, mask, g = y.cuda(), mask.cuda(), g.cuda()
        with torch.no_grad():
            want, _ = dec(y, mask, g=g, reverse=True)
        rev = GraphedReverse(dec, torch.zeros_like(y), mask, g=torch.zeros_like(g))
        got = rev(y, mask, g).clone()
        assert rev.launches > 0
        assert torch.equal(got, want)
        y2 = torch.randn_like(y) * mask
        with torch.no_grad():
            want2, _ = dec(y2, mask, g=g, reverse=True)
        assert torch.equal(rev(y2).clone(), want2)


@pytest.mark.parametrize("precision,segments", [("bf16", 3), ("bf16", 5), ("fp32", 2)])
def test_segmented_backward_equals_one_call_backward(precision, segments):
    """gtk_decoder_backward_range: the backward issued as consecutive block ranges (what the data-parallel path does to
    overlap the gradient all-reduce) produces the gradients of the single call."""
    import helpers
    from glowtts_b200 import models, ops
    cfg = dict(helpers.TINY_CFG, in_channels=16, n_blocks=5, n_layers=2, hidden_channels=64)
    sd = helpers.make_state_dict(cfg, 71)
    dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                 cfg["n_blocks"], cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"], precision=precision)
    dec.load_state_dict(sd, strict=True)
    dec = dec.cuda().train()
    y, mask, _, _ = helpers.flow_inputs(cfg, 3, 42, 5, False)
    y, mask = y.cuda(), mask.cuda()
    params, folded = dec._params()
    flat, offsets = dec._ensure_flat(params)
    mask2 = mask.reshape(mask.shape[0], mask.shape[-1])
    c = dec._cfg(folded)
    with torch.no_grad():
        z, logdet, ws = ops.decoder_forward_flat(y, mask2, flat, offsets, None, c, True, None)
        gen = torch.Generator(device="cuda").manual_seed(3)
        dz = torch.randn(z.shape, device="cuda", generator=gen) * mask
        dld = torch.randn(logdet.shape, device="cuda", generator=gen)
        dx0, df0, *_ = ops.decoder_backward_flat(dz, dld, mask2, flat, offsets, None, ws, c, y.shape[2], None)
        dx1, df1, *_ = ops.decoder_backward_flat_overlapped(dz, dld, mask2, flat, offsets, None, ws, c, y.shape[2], None, None, segments)
    assert torch.allclose(dx0, dx1, rtol=1e-5, atol=1e-6)
    assert torch.allclose(df0, df1, rtol=1e-5, atol=1e-6)
    assert df0.abs().sum().item() > 0


def test_graphed_step_prefetch_feeds_the_same_batches():
    """GraphedTrainStep.prefetch (copy-stream double buffering of the pinned host batch) == step(batch)."""
    import helpers
    from glowtts_b200 import models
    from glowtts_b200.hot_path import GraphedTrainStep
    cfg = dict(helpers.TINY_CFG, in_channels=16, n_blocks=2, n_layers=2, hidden_channels=64)
    sd = helpers.make_state_dict(cfg, 81)
    gen = torch.Generator().manual_seed(4)
    batches = []
    for _ in range(3):
        y = torch.randn(3, 16, 40, generator=gen)
        batches.append([y.pin_memory(), torch.tensor([40, 22, 30]).pin_memory(),
                        (torch.randn(3, 16, 20, generator=gen) * 0.5).pin_memory(), torch.tensor([20, 11, 15]).pin_memory()])
    losses = []
    for mode in ("direct", "prefetch"):
        dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                     cfg["n_blocks"], cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"], precision="bf16")
        dec.load_state_dict(sd, strict=True)
        dec = dec.cuda().train()
        opt = torch.optim.AdamW(dec.flat_parameters(), lr=1e-3, fused=True, capturable=True)
        st = GraphedTrainStep(dec, opt, [t.cuda() for t in batches[0]], warmup=1)
        ls = []
        if mode == "direct":
            for bt in batches:
                ls.append(st.step(bt).item())
        else:
            st.prefetch(batches[0])
            for i in range(3):
                loss = st.step()
                if i + 1 < 3:
                    st.prefetch(batches[i + 1])
                ls.append(loss.item())
        losses.append(ls)
    # same batches in the same order -> same losses up to the atomics' summation order in the backward
    assert all(abs(a - b) < 2e-4 * max(1.0, abs(a)) for a, b in zip(*losses)), losses
    assert abs(losses[0][0] - losses[0][1]) > 1e-3, losses      # the batches do differ: a stale buffer would show


def test_pdl_on_off_agree(tmp_path):
    """Programmatic dependent launch must not change a single forward value: the decoder at the BASELINE batch shape with
    GTK_PDL=1 (kernels start while their predecessor drains) and GTK_PDL=0 (plain stream order), three repetitions each.
    Forward / reverse outputs are deterministic -> bit-identical; gradients go through atomics -> 1e-5 relative."""
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    res = {}
    for v in ("1", "0"):
        path = str(tmp_path / f"pdl{v}.npz")
        env = dict(os.environ, GTK_PDL=v)
        subprocess.run([sys.executable, os.path.join(here, "pdl_worker.py"), path], check=True, env=env, timeout=300)
        res[v] = np.load(path)
    on, off = res["1"], res["0"]
    for rep in range(3):
        for k in ("z", "ld", "yb"):
            # ld accumulates per-utterance sums with atomics; z / yb are deterministic
            a, b = on[f"{k}{rep}"], off[f"{k}{rep}"]
            if k == "ld":
                np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-3)
            else:
                assert np.array_equal(a, b), (k, rep, np.abs(a - b).max())
        np.testing.assert_allclose(on[f"dy{rep}"], off[f"dy{rep}"], rtol=1e-4, atol=1e-7)
        ga, gb = on[f"g{rep}"], off[f"g{rep}"]
        assert np.linalg.norm(ga - gb) <= 1e-5 * np.linalg.norm(gb)
        assert np.array_equal(on[f"z{rep}"], on["z0"])
