"""oracle/flow.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU fp32 (torch) restatement of the reference's flow decoder, logp construction
and the glue either side of it, written as pure functions over a ``state_dict``
so that the same checkpoint drives the reference modules, this oracle and the
CUDA path.  Each function cites the reference lines it follows (paths relative
to /root/reference).  tests/test_oracle_flow.py pins these functions against the
imported reference modules (when /root/reference is present) and against the
committed fixtures in tests/golden/ (generated from the reference modules by
tests/golden/make_golden.py).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
SD = Dict[str, Tensor]


# ----------------------------------------------------------------------------- glue
def sequence_mask(length: Tensor, max_length: Optional[int] = None) -> Tensor:
    """commons.py:82-86."""
    if max_length is None:
        max_length = int(length.max())
    pos = torch.arange(max_length, dtype=length.dtype, device=length.device)
    return pos[None, :] < length[:, None]


def squeeze(x: Tensor, x_mask: Optional[Tensor] = None, n_sqz: int = 2):
    """commons.py:339-351: x[b,c,n*t'+s] -> out[b, s*C+c, t'], mask taken at the LAST
    frame of each group (x_mask[:,:,n_sqz-1::n_sqz]) and applied to the output."""
    b, c, t = x.shape
    tq = t // n_sqz
    x = x[:, :, : tq * n_sqz]
    out = x.reshape(b, c, tq, n_sqz).permute(0, 3, 1, 2).reshape(b, c * n_sqz, tq)
    if x_mask is not None:
        m = x_mask[:, :, n_sqz - 1 :: n_sqz]
    else:
        m = torch.ones(b, 1, tq, dtype=x.dtype, device=x.device)
    return out * m, m


def unsqueeze(x: Tensor, x_mask: Optional[Tensor] = None, n_sqz: int = 2):
    """commons.py:354-364: inverse index map of squeeze; mask repeated n_sqz times."""
    b, c, t = x.shape
    out = x.reshape(b, n_sqz, c // n_sqz, t).permute(0, 2, 3, 1).reshape(b, c // n_sqz, t * n_sqz)
    if x_mask is not None:
        m = x_mask.unsqueeze(-1).expand(-1, -1, -1, n_sqz).reshape(b, 1, t * n_sqz)
    else:
        m = torch.ones(b, 1, t * n_sqz, dtype=x.dtype, device=x.device)
    return out * m, m


def fused_add_tanh_sigmoid_multiply(a: Tensor, b: Tensor, n_channels: int) -> Tensor:
    """commons.py:61-68."""
    s = a + b
    return torch.tanh(s[:, :n_channels]) * torch.sigmoid(s[:, n_channels:])


def mle_loss(z: Tensor, m: Tensor, logs: Tensor, logdet: Tensor, mask: Tensor) -> Tensor:
    """commons.py:28-33 (first two sums deliberately unmasked, as in the reference)."""
    nll = logs.sum() + 0.5 * (torch.exp(-2.0 * logs) * (z - m) ** 2).sum()
    nll = nll - logdet.sum()
    nll = nll / (torch.ones_like(z) * mask).sum()
    return nll + 0.5 * math.log(2.0 * math.pi)


def logp(x_m: Tensor, x_logs: Tensor, z: Tensor) -> Tensor:
    """models.py:1077-1082: log N(z[:, :, y]; x_m[:, :, x], exp(x_logs[:, :, x])) as
    four terms, two of them [B,T_x,80]x[B,80,T_y] matmuls."""
    s = torch.exp(-2.0 * x_logs)
    t1 = (-0.5 * math.log(2.0 * math.pi) - x_logs).sum(1).unsqueeze(-1)
    t2 = torch.matmul(s.transpose(1, 2), -0.5 * z * z)
    t3 = torch.matmul((x_m * s).transpose(1, 2), z)
    t4 = (-0.5 * x_m * x_m * s).sum(1).unsqueeze(-1)
    return t1 + t2 + t3 + t4


def clip_grad_value_(grads, clip_value, norm_type=2):
    """commons.py:320-336 on a list of gradient tensors (modified in place) -> total norm (python float)."""
    norm_type = float(norm_type)
    total = 0.0
    for g in grads:
        total += float(g.norm(norm_type)) ** norm_type
        if clip_value is not None:
            g.clamp_(min=-float(clip_value), max=float(clip_value))
    return total ** (1.0 / norm_type)


def generate_path(duration: Tensor, mask: Tensor) -> Tensor:
    """commons.py:127-143: durations [B,T_x] -> hard alignment [B,T_x,T_y]."""
    b, t_x, t_y = mask.shape
    cum = torch.cumsum(duration, 1)
    upto = sequence_mask(cum.reshape(b * t_x), t_y).to(mask.dtype).reshape(b, t_x, t_y)
    prev = F.pad(upto, (0, 0, 1, 0))[:, :-1]
    return (upto - prev) * mask


def expand_prior(attn: Tensor, x_m: Tensor, x_logs: Tensor):
    """models.py:1118-1119: z_m[b,:,y] = x_m[b,:,x(y)] written as attn^T matmuls."""
    z_m = torch.matmul(attn.transpose(1, 2), x_m.transpose(1, 2)).transpose(1, 2)
    z_logs = torch.matmul(attn.transpose(1, 2), x_logs.transpose(1, 2)).transpose(1, 2)
    return z_m, z_logs


# ----------------------------------------------------------------------------- params
def effective_weight(sd: SD, prefix: str) -> Tensor:
    """weight of a (possibly weight-normed) conv: g * v / ||v|| with the norm over
    every dim but 0 (torch.nn.utils.weight_norm, used at modules.py:124,131,141 and
    attentions.py:103); plain ``weight`` after remove_weight_norm (modules.py:173-179)."""
    if prefix + ".weight" in sd:
        return sd[prefix + ".weight"]
    g, v = sd[prefix + ".weight_g"], sd[prefix + ".weight_v"]
    norm = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
    return v * (g / norm)


def _conv(sd: SD, prefix: str, x: Tensor, dilation: int = 1) -> Tensor:
    w = effective_weight(sd, prefix)
    k = w.shape[-1]
    pad = (k * dilation - dilation) // 2
    return F.conv1d(x, w, sd.get(prefix + ".bias"), padding=pad, dilation=dilation)


# ----------------------------------------------------------------------------- flows
def actnorm(x: Tensor, x_mask: Tensor, logs: Tensor, bias: Tensor, reverse: bool = False):
    """modules.py:584-599."""
    x_len = x_mask.sum([1, 2])
    if reverse:
        return (x - bias) * torch.exp(-logs) * x_mask, None
    return (bias + torch.exp(logs) * x) * x_mask, logs.sum() * x_len


def actnorm_ddi(x: Tensor, x_mask: Tensor):
    """modules.py:607-619: data-dependent init -> (logs, bias) as [1,C,1]."""
    denom = x_mask.sum([0, 2])
    mean = (x * x_mask).sum([0, 2]) / denom
    var = (x * x * x_mask).sum([0, 2]) / denom - mean**2
    half_logvar = 0.5 * torch.log(var.clamp_min(1e-6))
    return (-half_logvar).reshape(1, -1, 1), (-mean * torch.exp(-half_logvar)).reshape(1, -1, 1)


def invconvnear(x: Tensor, x_mask: Tensor, weight: Tensor, reverse: bool = False,
                n_split: int = 4, weight_inv: Optional[Tensor] = None):
    """modules.py:635-665 stated with the explicit index map of SURVEY 8a-5:
    channel c = h*(C/2) + g*(n_split/2) + k  (h in {0,1}, k < n_split/2) is entry
    j = h*(n_split/2) + k of group g's n_split-vector; z_j' = sum_j W[j',j] x_j."""
    b, c, t = x.shape
    half = n_split // 2
    groups = c // n_split
    x_len = x_mask.sum([1, 2])
    if reverse:
        w = weight_inv if weight_inv is not None else torch.inverse(weight.float()).to(weight.dtype)
        logdet = None
    else:
        w = weight
        logdet = torch.logdet(weight) * (c / n_split) * x_len
    v = x.reshape(b, 2, groups, half, t)                   # [b, h, g, k, t]
    z = torch.einsum("phqk,bhgkt->bpgqt", w.reshape(2, half, 2, half).permute(0, 2, 1, 3), v)
    # einsum indices: p=h', q=k' (output), h,k (input); result [b, h', g, k', t]
    z = z.reshape(b, c, t) * x_mask
    return z, logdet


def wn(sd: SD, prefix: str, x: Tensor, x_mask: Tensor, g: Optional[Tensor], hidden: int,
       dilation_rate: int, n_layers: int, drop=None) -> Tensor:
    """modules.py:144-171.  `drop(layer, x_in) -> x_in` stands for self.drop (modules.py:153); None = eval / p=0.
    The tests pass the keep-mask of the implementation under test, since nn.Dropout's RNG stream is not part of
    the contract."""
    out = torch.zeros_like(x)
    if g is not None:
        g = _conv(sd, prefix + ".cond_layer", g)
    for i in range(n_layers):
        x_in = _conv(sd, f"{prefix}.in_layers.{i}", x, dilation=dilation_rate**i)
        if drop is not None:
            x_in = drop(i, x_in)
        if g is not None:
            g_l = g[:, 2 * hidden * i : 2 * hidden * (i + 1)]
        else:
            g_l = torch.zeros_like(x_in)
        acts = fused_add_tanh_sigmoid_multiply(x_in, g_l, hidden)
        rs = _conv(sd, f"{prefix}.res_skip_layers.{i}", acts)
        if i < n_layers - 1:
            x = (x + rs[:, :hidden]) * x_mask
            out = out + rs[:, hidden:]
        else:
            out = out + rs
    return out * x_mask


def wnp(sd: SD, prefix: str, x: Tensor, x_mask: Tensor, g1: Optional[Tensor], hidden: int,
        dilation_rate: int, n_layers: int, n_sqz: int = 2, drop=None) -> Tensor:
    """modules.py:316-343 + 353-362: identity when g1 is None; otherwise the WN body
    with a time-varying conditioner cond_layer1(g1) squeezed by n_sqz.  `drop(layer, x_in)` stands for self.drop
    (modules.py:331), as in wn()."""
    if g1 is None:
        return x
    cond = _conv(sd, prefix + ".cond_layer1", g1)
    if n_sqz > 1:
        cond, _ = squeeze(cond, None, n_sqz)
    out = torch.zeros_like(x)
    for i in range(n_layers):
        x_in = _conv(sd, f"{prefix}.in_layers.{i}", x, dilation=dilation_rate**i)
        if drop is not None:
            x_in = drop(i, x_in)
        acts = fused_add_tanh_sigmoid_multiply(x_in, cond[:, 2 * hidden * i : 2 * hidden * (i + 1)], hidden)
        rs = _conv(sd, f"{prefix}.res_skip_layers.{i}", acts)
        if i < n_layers - 1:
            x = (x + rs[:, :hidden]) * x_mask
            out = out + rs[:, hidden:]
        else:
            out = out + rs
    return out * x_mask


def coupling(sd: SD, prefix: str, x: Tensor, x_mask: Tensor, *, hidden: int, dilation_rate: int,
             n_layers: int, g: Optional[Tensor] = None, pitch: Optional[Tensor] = None,
             energy: Optional[Tensor] = None, reverse: bool = False, sigmoid_scale: bool = False,
             n_sqz: int = 2, drop=None):
    """attentions.py:132-186."""
    c = x.shape[1]
    x0, x1 = x[:, : c // 2], x[:, c // 2 :]
    if pitch is not None and pitch.dim() == 2:
        pitch = pitch.unsqueeze(1)
    if energy is not None and energy.dim() == 2:
        energy = energy.unsqueeze(1)
    h = _conv(sd, prefix + ".start", x0) * x_mask
    h = wn(sd, prefix + ".wn", h, x_mask, g, hidden, dilation_rate, n_layers, drop)
    # the drop hook numbers the layers of the three WaveNets 0..L-1 (wn), L..2L-1 (wn_energy), 2L..3L-1 (wn_pitch)
    sub = lambda k: None if drop is None else (lambda i, v: drop(k * n_layers + i, v))
    h = wnp(sd, prefix + ".wn_energy", h, x_mask, energy, hidden, dilation_rate, n_layers, n_sqz, sub(1))
    h = wnp(sd, prefix + ".wn_pitch", h, x_mask, pitch, hidden, dilation_rate, n_layers, n_sqz, sub(2))
    out = _conv(sd, prefix + ".end", h)
    m, logs = out[:, : c // 2], out[:, c // 2 :]
    if sigmoid_scale:
        logs = torch.log(1e-6 + torch.sigmoid(logs + 2))
    if reverse:
        z1 = (x1 - m) * torch.exp(-logs) * x_mask
        logdet = None
    else:
        z1 = (m + torch.exp(logs) * x1) * x_mask
        logdet = (logs * x_mask).sum([1, 2])
    return torch.cat([x0, z1], 1), logdet


def decoder(sd: SD, x: Tensor, x_mask: Tensor, *, hidden: int = 192, dilation_rate: int = 1,
            n_blocks: int = 12, n_layers: int = 4, n_split: int = 4, n_sqz: int = 2,
            sigmoid_scale: bool = False, g: Optional[Tensor] = None, pitch: Optional[Tensor] = None,
            energy: Optional[Tensor] = None, reverse: bool = False, prefix: str = "", drop=None):
    """models.py:765-785: squeeze -> (ActNorm, InvConvNear, CouplingBlock) x n_blocks ->
    unsqueeze; reversed order with inverse maps and no logdet when reverse=True."""
    if n_sqz > 1:
        x, x_mask = squeeze(x, x_mask, n_sqz)
    logdet_tot = None if reverse else 0
    order = range(n_blocks) if not reverse else reversed(range(n_blocks))
    for blk in order:
        an, ic, cb = (f"{prefix}flows.{3 * blk + j}" for j in range(3))
        steps = [
            lambda t: actnorm(t, x_mask, sd[an + ".logs"], sd[an + ".bias"], reverse),
            lambda t: invconvnear(t, x_mask, sd[ic + ".weight"], reverse, n_split),
            lambda t: coupling(sd, cb, t, x_mask, hidden=hidden, dilation_rate=dilation_rate,
                               n_layers=n_layers, g=g, pitch=pitch, energy=energy, reverse=reverse,
                               sigmoid_scale=sigmoid_scale, n_sqz=n_sqz,
                               drop=None if drop is None else (lambda i, v, _b=blk: drop(_b, i, v))),
        ]
        if reverse:
            steps.reverse()
        for step in steps:
            x, ld = step(x)
            if not reverse:
                logdet_tot = logdet_tot + ld
    if n_sqz > 1:
        x, x_mask = unsqueeze(x, x_mask, n_sqz)
    return x, logdet_tot
