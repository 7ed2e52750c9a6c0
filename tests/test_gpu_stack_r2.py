"""The whole-WaveNet kernel (csrc/wn_stack.cu: start conv + every layer in one launch, 128-row windows with a recomputed
halo) against the layer-by-layer kernels (csrc/wn_tc.cu) it replaces: same epilogue arithmetic, so forward outputs, saved
activations and the data gradient must agree BIT FOR BIT (sums accumulated with atomics: to 1e-5).  The reference-facing parity of either path is pinned by
test_gpu_parity_r2.py / test_gpu_prosody_r2.py (which now run the stack kernel); this file pins the two against each other
on ragged multi-window batches (window seams, packing gaps, dead windows), with speaker conditioning, dropout, dilation
and the chained pitch / energy WaveNets."""
import os

import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu

BASE = dict(in_channels=80, hidden_channels=192, kernel_size=5, dilation_rate=1, n_blocks=2, n_layers=4, n_split=4,
            n_sqz=2, gin_channels=0)
BASE_G = dict(BASE, gin_channels=16)
H128_DIL = dict(in_channels=16, hidden_channels=128, kernel_size=3, dilation_rate=2, n_blocks=2, n_layers=4, n_split=4,
                n_sqz=2, gin_channels=8)
H64_PE = dict(in_channels=16, hidden_channels=64, kernel_size=3, dilation_rate=2, n_blocks=2, n_layers=3, n_split=4,
              n_sqz=2, gin_channels=8)


def _decoder(cfg, sd, p_dropout):
    from glowtts_b200 import models
    dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                 cfg["n_blocks"], cfg["n_layers"], p_dropout=p_dropout, n_split=cfg["n_split"],
                                 n_sqz=cfg["n_sqz"], gin_channels=cfg["gin_channels"], precision="bf16")
    dec.load_state_dict(sd, strict=True)
    return dec.cuda()


def _run(cfg, sd, y, mask, cond, pitch, energy, p_dropout, stack):
    from glowtts_b200 import commons, _lib
    os.environ["GTK_WN_STACK"] = "1" if stack else "0"
    try:
        torch.manual_seed(1234)                           # the per-call dropout seed starts from torch's CPU generator
        dec = _decoder(cfg, sd, p_dropout)
        dec.train()
        yy = y.clone().requires_grad_(True)
        kw = {}
        if pitch is not None:
            kw["pitch"] = pitch.clone().requires_grad_(True)
        if energy is not None:
            kw["energy"] = energy.clone().requires_grad_(True)
        n0 = _lib.launch_count()
        z, ld = dec(yy, mask, g=cond, **kw)
        n_fwd = _lib.launch_count() - n0
        loss = commons.mle_loss(z, None, None, ld, mask)
        loss.backward()
        with torch.no_grad():
            dec.eval()
            y_rev, _ = dec(z.detach(), mask, g=cond, reverse=True, **{k: v.detach() for k, v in kw.items()})
        grads = {k: p.grad.clone() for k, p in dec.named_parameters() if p.grad is not None}
        extra = {k: v.grad.clone() for k, v in kw.items()}
        return dict(z=z.detach(), ld=ld.detach(), dy=yy.grad.clone(), y_rev=y_rev, grads=grads, extra=extra, n_fwd=n_fwd)
    finally:
        os.environ.pop("GTK_WN_STACK", None)


ONE_LAYER = dict(in_channels=16, hidden_channels=64, kernel_size=5, dilation_rate=1, n_blocks=2, n_layers=1, n_split=4,
                 n_sqz=2, gin_channels=0)
DIL3 = dict(in_channels=32, hidden_channels=128, kernel_size=3, dilation_rate=3, n_blocks=1, n_layers=3, n_split=4,
            n_sqz=2, gin_channels=0)       # halos 1, 3, 9: windows own 102 of 128 rows, 9 rows of slack either side

CASES = [
    ("base", BASE, [860, 430, 612, 2, 256, 700], 0.0, False),
    ("base_cond_dropout", BASE_G, [300, 226, 1024, 58], 0.05, False),
    ("h128_dilated", H128_DIL, [700, 514, 2, 388, 130], 0.0, False),
    ("h64_pitch_energy", H64_PE, [700, 514, 2, 388, 130], 0.05, True),
    ("one_layer", ONE_LAYER, [300, 2, 258, 64], 0.0, False),
    ("dilation_3", DIL3, [612, 10, 430], 0.05, False),
]


@pytest.mark.parametrize("name,cfg,lengths,p_dropout,prosody", CASES, ids=[c[0] for c in CASES])
def test_stack_kernel_is_bit_identical_to_layer_kernels(name, cfg, lengths, p_dropout, prosody):
    sd = helpers.make_state_dict(cfg, 71)
    B, T = len(lengths), max(lengths)
    gen = torch.Generator().manual_seed(72)
    lens = torch.tensor(lengths)
    mask = (torch.arange(T)[None, :] < lens[:, None]).float().unsqueeze(1).cuda()
    y = torch.randn(B, cfg["in_channels"], T, generator=gen).cuda() * mask
    cond = torch.randn(B, cfg["gin_channels"], 1, generator=gen).cuda() if cfg["gin_channels"] else None
    pitch = torch.randn(B, 1, T, generator=gen).cuda() if prosody else None
    energy = torch.randn(B, T, generator=gen).cuda() if prosody else None
    a = _run(cfg, sd, y, mask, cond, pitch, energy, p_dropout, stack=True)
    b = _run(cfg, sd, y, mask, cond, pitch, energy, p_dropout, stack=False)
    n_wn = 3 if prosody else 1
    # the stack path is one launch per WaveNet instead of start + L layers, and `end` + the affine coupling ride in the
    # last WaveNet's launch
    assert b["n_fwd"] - a["n_fwd"] == cfg["n_blocks"] * (cfg["n_layers"] * n_wn + 2 - n_wn), (a["n_fwd"], b["n_fwd"])
    assert torch.isfinite(a["z"]).all()
    for k in ("z", "y_rev", "dy"):
        assert torch.equal(a[k], b[k]), (name, k, float((a[k] - b[k]).abs().max()))
    for k in b["extra"]:
        assert torch.equal(a["extra"][k], b["extra"][k]), (name, k)
    # sums that either path accumulates with atomicAdd (per-utterance logdet, bias / ActNorm / 4x4-mix gradients) differ
    # from run to run in their last bits
    assert torch.allclose(a["ld"], b["ld"], rtol=1e-5, atol=1e-4), (a["ld"], b["ld"])
    assert set(a["grads"]) == set(b["grads"])
    for k in b["grads"]:
        ga, gb = a["grads"][k].double(), b["grads"][k].double()
        assert float((ga - gb).norm()) <= 1e-5 * float(gb.norm()) + 1e-12, (name, k, float((ga - gb).abs().max()))


@pytest.mark.parametrize("name,cfg,lengths,p_dropout,prosody", [CASES[0], CASES[3], CASES[5]],
                         ids=[CASES[0][0], CASES[3][0], CASES[5][0]])
def test_start_data_gradient_inside_the_layer_0_backward_is_bit_identical(name, cfg, lengths, p_dropout, prosody):
    """The data gradient of the `start` 1x1 conv rides in the layer-0 launch of csrc/wn_bwd_tc.cu (GEMM c on the d(h_0) tile
    in shared memory, same two-accumulator sums as csrc/conv_tc.cu): one launch fewer per coupling block and the same bits
    as the separate conv (GTK_WB_POST=0)."""
    from glowtts_b200 import _lib
    sd = helpers.make_state_dict(cfg, 81)
    B, T = len(lengths), max(lengths)
    gen = torch.Generator().manual_seed(82)
    lens = torch.tensor(lengths)
    mask = (torch.arange(T)[None, :] < lens[:, None]).float().unsqueeze(1).cuda()
    y = torch.randn(B, cfg["in_channels"], T, generator=gen).cuda() * mask
    cond = torch.randn(B, cfg["gin_channels"], 1, generator=gen).cuda() if cfg["gin_channels"] else None
    pitch = torch.randn(B, 1, T, generator=gen).cuda() if prosody else None
    energy = torch.randn(B, T, generator=gen).cuda() if prosody else None
    out = {}
    for post in ("1", "0"):
        os.environ["GTK_WB_POST"] = post
        try:
            n0 = _lib.launch_count()
            out[post] = _run(cfg, sd, y, mask, cond, pitch, energy, p_dropout, stack=True)
            out[post]["n"] = _lib.launch_count() - n0
        finally:
            os.environ.pop("GTK_WB_POST", None)
    assert out["0"]["n"] - out["1"]["n"] == cfg["n_blocks"], (out["0"]["n"], out["1"]["n"])
    assert torch.isfinite(out["1"]["dy"]).all()
    assert torch.equal(out["1"]["dy"], out["0"]["dy"]), float((out["1"]["dy"] - out["0"]["dy"]).abs().max())
    for k in out["0"]["extra"]:
        assert torch.equal(out["1"]["extra"][k], out["0"]["extra"][k]), k
    for k in out["0"]["grads"]:
        ga, gb = out["1"]["grads"][k].double(), out["0"]["grads"][k].double()
        assert float((ga - gb).norm()) <= 1e-5 * float(gb.norm()) + 1e-12, (name, k)
