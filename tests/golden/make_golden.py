"""Generate tests/golden/*.npz by running the REFERENCE's own code in this container.

  python tests/golden/make_golden.py

* mas_*.npz     : seeded value/lengths -> alignment from the reference's compiled
                  monotonic_align/core.pyx (oracle/_ref, built by oracle/build.py).
* flow_*.npz    : seeded checkpoint (tests/helpers.make_state_dict) + inputs through the
                  imported reference modules (models.FlowSpecDecoder, modules.ActNorm /
                  InvConvNear / WN, attentions.CouplingBlock, commons.*) -> outputs/grads.
* logp_*.npz    : models.py:1077-1082 executed verbatim via exec of those source lines.

/root/reference cannot travel to the GPU box, hence committed fixtures.  Fixtures store
seeds + outputs (inputs are regenerated from the seed by tests/helpers), so they stay small.
"""
import math
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers  # noqa: E402
import refimport  # noqa: E402
from oracle import mas as omas  # noqa: E402

MAS_CASES = {
    # name: (b, t_x, t_y, seed, ragged, quantised)
    "mas_small_ragged": (4, 23, 57, 11, True, False),
    "mas_ties": (4, 19, 40, 12, True, True),
    "mas_square": (2, 16, 16, 13, False, False),       # t_x == t_y: forced diagonal
    "mas_tx1": (2, 1, 9, 14, False, False),            # t_x == 1
    "mas_medium": (3, 70, 150, 15, True, False),
    "mas_wide": (2, 97, 333, 16, True, True),          # > 3 warps of x, ties
}


def gen_mas():
    for name, (b, tx, ty, seed, ragged, quant) in MAS_CASES.items():
        value, mask, t_xs, t_ys = helpers.mas_case(b, tx, ty, seed, ragged, quant)
        path = omas.maximum_path_np(value, mask, impl="ref")
        np.savez_compressed(os.path.join(HERE, name + ".npz"), case=np.array([b, tx, ty, seed, ragged, quant]),
                            t_xs=t_xs, t_ys=t_ys, index=helpers.path_to_index(path),
                            path_sum=path.sum((1, 2)).astype(np.int32))
        print(name, path.sum())


FLOW_CASES = {
    # name: (cfg, sd_seed, b, t, in_seed, use_g)
    "flow_tiny": (helpers.TINY_CFG, 21, 3, 14, 31, False),
    "flow_tiny_g": (dict(helpers.TINY_CFG, gin_channels=6), 22, 2, 12, 32, True),
    "flow_tiny_k3_dil2": (dict(helpers.TINY_CFG, kernel_size=3, dilation_rate=2, n_layers=3), 23, 2, 20, 33, False),
    "flow_base": (helpers.BASE_CFG, 24, 2, 36, 34, False),
}
END_STD = {"flow_base": 0.01}  # 12 blocks x hidden 192: keep exp(logs) tame so the flow stays invertible in fp32


flow_inputs = helpers.flow_inputs


def gen_flow(ref):
    for name, (cfg, sd_seed, b, t, in_seed, use_g) in FLOW_CASES.items():
        dec = ref.models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"],
                                         cfg["dilation_rate"], cfg["n_blocks"], cfg["n_layers"], p_dropout=0.0,
                                         n_split=cfg["n_split"], n_sqz=cfg["n_sqz"], gin_channels=cfg["gin_channels"])
        sd = helpers.make_state_dict(cfg, sd_seed, end_std=END_STD.get(name, 0.05))
        dec.load_state_dict(sd, strict=True)
        dec.eval()
        y, mask, lengths, g = flow_inputs(cfg, b, t, in_seed, use_g)
        y = y.clone().requires_grad_(True)
        z, logdet = dec(y, mask, g=g, reverse=False)
        # mle-style scalar so that grads exercise both z and logdet (commons.py:28-33)
        zm = torch.zeros_like(z)
        loss = ref.commons.mle_loss(z, zm, torch.zeros_like(z), logdet, mask)
        loss.backward()
        grads = {k: p.grad.detach().numpy() for k, p in dec.named_parameters() if p.grad is not None}
        # a handful of parameter grads is enough to pin the backward (keeps the file small)
        keep = [k for k in grads if (".wn_pitch." not in k and ".wn_energy." not in k)]
        sel = {}
        for k in keep:
            a = grads[k]
            if cfg is helpers.BASE_CFG or cfg.get("hidden_channels") == 192:
                blk = int(k.split(".")[1]) // 3
                if blk not in (0, 11):
                    continue
                if a.size > 4096:
                    a = a.reshape(-1)[:: max(1, a.size // 2048)]  # strided sample
            sel["grad:" + k] = a
        with torch.no_grad():
            y_rev, _ = dec(z.detach(), mask, g=g, reverse=True)
            dec.store_inverse()
            y_rev2, _ = dec(z.detach(), mask, g=g, reverse=True)
        np.savez_compressed(os.path.join(HERE, name + ".npz"),
                            sd_seed=sd_seed, b=b, t=t, in_seed=in_seed, use_g=use_g, end_std=END_STD.get(name, 0.05),
                            z=z.detach().numpy(), logdet=logdet.detach().numpy(), loss=loss.item(),
                            dy=y.grad.numpy(), y_rev=y_rev.numpy(), y_rev_stored=y_rev2.numpy(), **sel)
        print(name, float(loss), float((y_rev - y.detach()).abs().max()))


PROSODY_CASES = {
    # name: (cfg, sd_seed, b, t, in_seed, use_g, use_pitch, use_energy)   -- the live fork's 3-WN coupling blocks
    "flow_tiny_pe": (dict(helpers.TINY_CFG, gin_channels=6), 25, 3, 14, 35, True, True, True),
    "flow_tiny_p": (helpers.TINY_CFG, 26, 2, 12, 36, False, True, False),
}
PROSODY_H64_CASES = {
    # round 2: tensor-core-capable widths (hidden 64) so that the fused 3-WN chain of the bf16 path has reference fixtures
    "flow_h64_pe": (dict(in_channels=16, hidden_channels=64, kernel_size=3, dilation_rate=2, n_blocks=2, n_layers=3,
                         n_split=4, n_sqz=2, gin_channels=8), 45, 3, 280, 46, True, True, True),
    "flow_h64_p": (dict(in_channels=16, hidden_channels=64, kernel_size=5, dilation_rate=1, n_blocks=2, n_layers=3,
                        n_split=4, n_sqz=2, gin_channels=0), 47, 2, 300, 48, False, True, False),
}


prosody_inputs = helpers.prosody_inputs


def gen_flow_prosody(ref, cases=None):
    """FlowSpecDecoder with pitch / energy tracks: every coupling block runs wn -> wn_energy -> wn_pitch
    (attentions.py:155-160, modules.WNP)."""
    for name, (cfg, sd_seed, b, t, in_seed, use_g, use_p, use_e) in (cases or PROSODY_CASES).items():
        dec = ref.models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"],
                                         cfg["dilation_rate"], cfg["n_blocks"], cfg["n_layers"], p_dropout=0.0,
                                         n_split=cfg["n_split"], n_sqz=cfg["n_sqz"], gin_channels=cfg["gin_channels"])
        sd = helpers.make_state_dict(cfg, sd_seed)
        dec.load_state_dict(sd, strict=True)
        dec.eval()
        y, mask, lengths, g = flow_inputs(cfg, b, t, in_seed, use_g)
        pitch, energy = prosody_inputs(b, t, in_seed, use_p, use_e)
        y = y.clone().requires_grad_(True)
        if pitch is not None:
            pitch = pitch.clone().requires_grad_(True)
        if energy is not None and cases is not None:
            energy = energy.clone().requires_grad_(True)
        z, logdet = dec(y, mask, g=g, pitch=pitch, energy=energy, reverse=False)
        loss = ref.commons.mle_loss(z, torch.zeros_like(z), torch.zeros_like(z), logdet, mask)
        loss.backward()
        sel = {}
        for k, p in dec.named_parameters():
            if p.grad is None:
                continue
            a = p.grad.detach().numpy()
            if cases is not None and a.size > 4096:
                a = a.reshape(-1)[:: max(1, a.size // 2048)]    # strided sample keeps the file small
            sel["grad:" + k] = a
        with torch.no_grad():
            y_rev, _ = dec(z.detach(), mask, g=g, pitch=pitch, energy=energy, reverse=True)
            dec.store_inverse()
            y_rev2, _ = dec(z.detach(), mask, g=g, pitch=pitch, energy=energy, reverse=True)
        extra = {} if pitch is None else {"dpitch": pitch.grad.numpy()}
        if energy is not None and energy.grad is not None:
            extra["denergy"] = energy.grad.numpy()
        np.savez_compressed(os.path.join(HERE, name + ".npz"),
                            sd_seed=sd_seed, b=b, t=t, in_seed=in_seed, use_g=use_g, use_pitch=use_p, use_energy=use_e,
                            z=z.detach().numpy(), logdet=logdet.detach().numpy(), loss=loss.item(),
                            dy=y.grad.numpy(), y_rev=y_rev.numpy(), y_rev_stored=y_rev2.numpy(), **extra, **sel)
        print(name, float(loss), float((y_rev - y.detach()).abs().max()), len(sel))


ROUND2_CASES = {
    # name: (cfg, sd_seed, b, t, in_seed, use_g, sigmoid_scale)  -- round 2: reference branches the round-1 fixtures missed
    "flow_tiny_sig": (helpers.TINY_CFG, 27, 3, 14, 37, False, True),                      # attentions.py:175-176
    "flow_h64": (dict(in_channels=16, hidden_channels=64, kernel_size=5, dilation_rate=1, n_blocks=3, n_layers=3,
                      n_split=4, n_sqz=2, gin_channels=0), 28, 3, 300, 38, False, False),   # runs on the bf16 tcgen05 path
    "flow_h64_g_sig": (dict(in_channels=16, hidden_channels=64, kernel_size=3, dilation_rate=2, n_blocks=2, n_layers=3,
                            n_split=4, n_sqz=2, gin_channels=8), 29, 2, 280, 39, True, True),
}


def gen_round2(ref):
    """Decoder fixtures with sigmoid_scale=True and tensor-core-capable widths (hidden 64), plus module fixtures with
    x_mask=None (modules.py:585-586, 639-641; attentions.py:134-135) and no_jacobian=True (modules.py:655-656)."""
    for name, (cfg, sd_seed, b, t, in_seed, use_g, sig) in ROUND2_CASES.items():
        dec = ref.models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"],
                                         cfg["dilation_rate"], cfg["n_blocks"], cfg["n_layers"], p_dropout=0.0,
                                         n_split=cfg["n_split"], n_sqz=cfg["n_sqz"], sigmoid_scale=sig,
                                         gin_channels=cfg["gin_channels"])
        sd = helpers.make_state_dict(cfg, sd_seed)
        dec.load_state_dict(sd, strict=True)
        dec.eval()
        y, mask, lengths, g = flow_inputs(cfg, b, t, in_seed, use_g)
        y = y.clone().requires_grad_(True)
        if g is not None:
            g = g.clone().requires_grad_(True)
        z, logdet = dec(y, mask, g=g, reverse=False)
        loss = ref.commons.mle_loss(z, torch.zeros_like(z), torch.zeros_like(z), logdet, mask)
        loss.backward()
        sel = {}
        for k, p in dec.named_parameters():
            if p.grad is None or ".wn_pitch." in k or ".wn_energy." in k:
                continue
            a = p.grad.detach().numpy()
            if a.size > 4096:
                a = a.reshape(-1)[:: max(1, a.size // 2048)]
            sel["grad:" + k] = a
        with torch.no_grad():
            y_rev, _ = dec(z.detach(), mask, g=g, reverse=True)
            dec.store_inverse()
            y_rev2, _ = dec(z.detach(), mask, g=g, reverse=True)
        extra = {} if g is None else {"dg": g.grad.numpy()}
        np.savez_compressed(os.path.join(HERE, name + ".npz"),
                            sd_seed=sd_seed, b=b, t=t, in_seed=in_seed, use_g=use_g, sigmoid_scale=sig, end_std=0.05,
                            z=z.detach().numpy().astype(np.float32), logdet=logdet.detach().numpy(), loss=loss.item(),
                            dy=y.grad.numpy(), y_rev=y_rev.numpy(), y_rev_stored=y_rev2.numpy(), **extra, **sel)
        print(name, float(loss), float((y_rev - y.detach()).abs().max()), len(sel))
    # ---- modules with x_mask=None / no_jacobian=True
    gen = torch.Generator().manual_seed(43)
    b, c, t, h = 2, 16, 9, 16
    x = torch.randn(b, c, t, generator=gen)
    out = {"x": x.numpy()}
    an = ref.modules.ActNorm(c)
    an.logs.data = 0.2 * torch.randn(1, c, 1, generator=gen)
    an.bias.data = 0.2 * torch.randn(1, c, 1, generator=gen)
    z, ld = an(x)
    zr, _ = an(x, reverse=True)
    out.update(an_logs=an.logs.data.numpy(), an_bias=an.bias.data.numpy(), an_z=z.detach().numpy(),
               an_logdet=ld.detach().numpy(), an_zrev=zr.detach().numpy())
    ic = ref.modules.InvConvNear(c, n_split=4)
    ic.weight.data = ic.weight.data + 0.1 * torch.randn(4, 4, generator=gen)
    z, ld = ic(x)
    zr, _ = ic(x, reverse=True)
    out.update(ic_w=ic.weight.data.numpy(), ic_z=z.detach().numpy(), ic_logdet=ld.detach().numpy(), ic_zrev=zr.detach().numpy())
    ic2 = ref.modules.InvConvNear(c, n_split=4, no_jacobian=True)
    ic2.weight.data = ic.weight.data.clone()
    lengths = torch.tensor([9, 5])
    mask = (torch.arange(t)[None, :] < lengths[:, None]).float().unsqueeze(1)
    z2, ld2 = ic2(x, mask)
    assert ld2 == 0
    out.update(nj_lengths=lengths.numpy(), nj_z=z2.detach().numpy())
    cfg = dict(helpers.TINY_CFG, n_blocks=1)
    for sig in (False, True):
        cb = ref.attentions.CouplingBlock(c, h, kernel_size=5, dilation_rate=1, n_layers=2, sigmoid_scale=sig)
        sd = {k[len("flows.2."):]: v for k, v in helpers.make_state_dict(cfg, 44).items() if k.startswith("flows.2.")}
        cb.load_state_dict(sd, strict=True)
        cb.eval()
        z, ld = cb(x)
        zr, _ = cb(x, reverse=True)
        tag = "cbs" if sig else "cb"
        out.update({tag + "_z": z.detach().numpy(), tag + "_logdet": ld.detach().numpy(), tag + "_zrev": zr.detach().numpy()})
    np.savez_compressed(os.path.join(HERE, "modules_nomask.npz"), **out)
    print("modules_nomask ok")


def gen_modules(ref):
    """Module-level fixtures: ActNorm (+DDI), InvConvNear, WN, CouplingBlock, gate, squeeze."""
    gen = torch.Generator().manual_seed(41)
    b, c, t, h = 3, 16, 11, 16
    x = torch.randn(b, c, t, generator=gen)
    lengths = torch.tensor([11, 7, 4])
    mask = (torch.arange(t)[None, :] < lengths[:, None]).float().unsqueeze(1)
    out = {"x": x.numpy(), "lengths": lengths.numpy()}
    an = ref.modules.ActNorm(c)
    an.logs.data = 0.2 * torch.randn(1, c, 1, generator=gen)
    an.bias.data = 0.2 * torch.randn(1, c, 1, generator=gen)
    z, ld = an(x, mask)
    zr, _ = an(x, mask, reverse=True)
    out.update(an_logs=an.logs.data.numpy(), an_bias=an.bias.data.numpy(), an_z=z.detach().numpy(),
               an_logdet=ld.detach().numpy(), an_zrev=zr.detach().numpy())
    an2 = ref.modules.ActNorm(c, ddi=True)
    z2, ld2 = an2(x, mask)
    out.update(ddi_logs=an2.logs.data.numpy(), ddi_bias=an2.bias.data.numpy(), ddi_z=z2.detach().numpy())
    ic = ref.modules.InvConvNear(c, n_split=4)
    ic.weight.data = ic.weight.data + 0.1 * torch.randn(4, 4, generator=gen)
    z, ld = ic(x, mask)
    zr, _ = ic(x, mask, reverse=True)
    out.update(ic_w=ic.weight.data.numpy(), ic_z=z.detach().numpy(), ic_logdet=ld.detach().numpy(),
               ic_zrev=zr.detach().numpy())
    a = torch.randn(b, 2 * h, t, generator=gen)
    bb = torch.randn(b, 2 * h, 1, generator=gen)
    out.update(gate_a=a.numpy(), gate_b=bb.numpy(),
               gate_out=ref.commons.fused_add_tanh_sigmoid_multiply(a, bb, torch.IntTensor([h])).numpy())
    xs, ms = ref.commons.squeeze(x, mask, 2)
    xu, mu = ref.commons.unsqueeze(xs, ms, 2)
    out.update(sq_x=xs.numpy(), sq_mask=ms.numpy(), unsq_x=xu.numpy(), unsq_mask=mu.numpy())
    dur = torch.tensor([[2., 3., 1., 0.], [1., 1., 4., 2.]])
    pm = torch.ones(2, 4, 9)
    pm[1, :, 8:] = 0
    out.update(gp_dur=dur.numpy(), gp_mask=pm.numpy(), gp_path=ref.commons.generate_path(dur, pm).numpy())
    np.savez_compressed(os.path.join(HERE, "modules_small.npz"), **out)
    print("modules_small ok")


def gen_logp():
    """Execute models.py:1077-1082 verbatim (read from the reference file at run time)."""
    src = open(os.path.join(refimport.REFERENCE_ROOT, "models.py")).read().splitlines()[1076:1082]
    code = "\n".join(line.strip() for line in src)
    assert code.startswith("x_s_sq_r") and "logp = logp1" in code, code
    gen = torch.Generator().manual_seed(51)
    b, d, tx, ty = 2, 80, 13, 29
    env = {"torch": torch, "math": math,
           "x_m": 0.5 * torch.randn(b, d, tx, generator=gen),
           "x_logs": 0.3 * torch.randn(b, d, tx, generator=gen),
           "z": torch.randn(b, d, ty, generator=gen)}
    exec(code, env)
    np.savez_compressed(os.path.join(HERE, "logp_small.npz"), x_m=env["x_m"].numpy(),
                        x_logs=env["x_logs"].numpy(), z=env["z"].numpy(), logp=env["logp"].numpy())
    print("logp_small ok")


if __name__ == "__main__":
    torch.manual_seed(0)
    ref = refimport.load()
    assert ref is not None, "needs /root/reference"
    if "--round2-only" in sys.argv:        # add the round-2 fixtures without touching the older files
        gen_round2(ref)
        sys.exit(0)
    if "--prosody-only" in sys.argv:       # add the 3-WN fixtures without touching the older files
        gen_flow_prosody(ref)
        sys.exit(0)
    if "--prosody-h64-only" in sys.argv:   # round 2: the hidden-64 3-WN fixtures
        gen_flow_prosody(ref, PROSODY_H64_CASES)
        sys.exit(0)
    gen_flow_prosody(ref)
    gen_flow_prosody(ref, PROSODY_H64_CASES)
    gen_round2(ref)
    gen_mas()
    gen_modules(ref)
    gen_logp()
    gen_flow(ref)
