"""Shared test helpers: deterministic decoder checkpoints and synthetic inputs.

``make_state_dict`` builds a checkpoint with exactly the key names / shapes the
reference's models.FlowSpecDecoder exposes (probed), from a
seed alone, so fixtures only need to store the seed -- the same function feeds the
reference modules (tests/golden/make_golden.py), the oracle and the CUDA path.
"""
from __future__ import annotations

from collections import OrderedDict

import numpy as np
import torch

BASE_CFG = dict(in_channels=80, hidden_channels=192, kernel_size=5, dilation_rate=1, n_blocks=12,
                n_layers=4, n_split=4, n_sqz=2, gin_channels=0)
TINY_CFG = dict(in_channels=8, hidden_channels=16, kernel_size=5, dilation_rate=1, n_blocks=2,
                n_layers=2, n_split=4, n_sqz=2, gin_channels=0)


def decoder_param_shapes(cfg) -> "OrderedDict[str, tuple]":
    c = cfg["in_channels"] * cfg["n_sqz"]
    h, k, nl, gin = cfg["hidden_channels"], cfg["kernel_size"], cfg["n_layers"], cfg["gin_channels"]
    out = OrderedDict()
    for b in range(cfg["n_blocks"]):
        an, ic, cb = (f"flows.{3 * b + j}" for j in range(3))
        out[an + ".logs"] = (1, c, 1)
        out[an + ".bias"] = (1, c, 1)
        out[ic + ".weight"] = (cfg["n_split"], cfg["n_split"])
        out[cb + ".start.bias"] = (h,)
        out[cb + ".start.weight_g"] = (h, 1, 1)
        out[cb + ".start.weight_v"] = (h, c // 2, 1)
        out[cb + ".end.weight"] = (c, h, 1)
        out[cb + ".end.bias"] = (c,)
        for wn_name in ("wn_pitch", "wn_energy", "wn"):
            p = f"{cb}.{wn_name}"
            for i in range(nl):
                out[f"{p}.in_layers.{i}.bias"] = (2 * h,)
                out[f"{p}.in_layers.{i}.weight_g"] = (2 * h, 1, 1)
                out[f"{p}.in_layers.{i}.weight_v"] = (2 * h, h, k)
            for i in range(nl):
                rs = 2 * h if i < nl - 1 else h
                out[f"{p}.res_skip_layers.{i}.bias"] = (rs,)
                out[f"{p}.res_skip_layers.{i}.weight_g"] = (rs, 1, 1)
                out[f"{p}.res_skip_layers.{i}.weight_v"] = (rs, h, 1)
            if wn_name == "wn":
                if gin:
                    out[p + ".cond_layer.bias"] = (2 * h * nl,)
                    out[p + ".cond_layer.weight_g"] = (2 * h * nl, 1, 1)
                    out[p + ".cond_layer.weight_v"] = (2 * h * nl, gin, 1)
            else:
                cc = 2 * h * nl // cfg["n_sqz"]
                out[p + ".cond_layer1.bias"] = (cc,)
                out[p + ".cond_layer1.weight_g"] = (cc, 1, 1)
                out[p + ".cond_layer1.weight_v"] = (cc, 1, 1)
    return out


def make_state_dict(cfg, seed: int, end_std: float = 0.05) -> "OrderedDict[str, torch.Tensor]":
    """Seeded, non-trivial checkpoint (couplings active, ActNorm non-identity, W general)."""
    gen = torch.Generator().manual_seed(seed)
    sd = OrderedDict()
    for name, shape in decoder_param_shapes(cfg).items():
        if name.endswith(".logs"):
            t = 0.1 * torch.randn(shape, generator=gen)
        elif name.endswith("weight_g"):
            t = 0.5 + torch.rand(shape, generator=gen)
        elif name.endswith("weight_v"):
            fan_in = int(np.prod(shape[1:]))
            t = torch.randn(shape, generator=gen) / np.sqrt(fan_in)
        elif ".end.weight" in name:
            t = end_std * torch.randn(shape, generator=gen)
        elif len(shape) == 2:  # InvConvNear 4x4: orthogonal + perturbation, det > 0
            q = torch.linalg.qr(torch.randn(shape, generator=gen))[0]
            q = q + 0.1 * torch.randn(shape, generator=gen)
            if torch.det(q) < 0:
                q[:, 0] = -q[:, 0]
            t = q
        else:  # biases
            t = 0.1 * torch.randn(shape, generator=gen)
        sd[name] = t.float().contiguous()
    return sd


def make_lengths(b: int, t_max: int, seed: int, lo_frac: float = 0.5, multiple: int = 1) -> torch.Tensor:
    gen = torch.Generator().manual_seed(seed)
    lo = max(multiple, int(t_max * lo_frac))
    ln = torch.randint(lo, t_max + 1, (b,), generator=gen)
    ln = (ln // multiple) * multiple
    ln[0] = (t_max // multiple) * multiple  # at least one full-length utterance
    return ln


def mas_case(b, t_x, t_y, seed, ragged=True, quantised=False):
    """Synthetic MAS input: value fp32 [b,t_x,t_y], rectangular prefix mask, lengths."""
    rng = np.random.default_rng(seed)
    value = rng.standard_normal((b, t_x, t_y)).astype(np.float32)
    if quantised:  # force exact ties
        value = (np.round(value * 4) / 4).astype(np.float32)
    if ragged:
        t_xs = rng.integers(max(1, t_x // 2), t_x + 1, b)
        t_ys = rng.integers(max(1, t_y // 2), t_y + 1, b)
        t_xs[0], t_ys[0] = t_x, t_y
        t_ys = np.maximum(t_ys, t_xs)
    else:
        t_xs = np.full(b, t_x)
        t_ys = np.full(b, t_y)
    mask = np.zeros((b, t_x, t_y), np.float32)
    for i in range(b):
        mask[i, : t_xs[i], : t_ys[i]] = 1
    return value, mask, t_xs.astype(np.int32), t_ys.astype(np.int32)


def path_to_index(path: np.ndarray) -> np.ndarray:
    """[b,t_x,t_y] 0/1 -> per-frame text index [b,t_y] (-1 where the column is empty)."""
    has = path.sum(1) > 0
    return np.where(has, path.argmax(1), -1).astype(np.int16)


def flow_inputs(cfg, b, t, seed, use_g):
    gen = torch.Generator().manual_seed(seed)
    y = torch.randn(b, cfg["in_channels"], t, generator=gen)
    lengths = make_lengths(b, t, seed + 1, multiple=cfg["n_sqz"])
    mask = (torch.arange(t)[None, :] < lengths[:, None]).float().unsqueeze(1)
    g = torch.randn(b, cfg["gin_channels"], 1, generator=gen) if use_g else None
    return y * mask, mask, lengths, g


def prosody_inputs(b, t, seed, use_pitch, use_energy):
    """Pitch [b,1,t] / energy [b,t] tracks of the live fork's coupling blocks (2-D on purpose: attentions.py:137-141)."""
    gen = torch.Generator().manual_seed(seed + 100)
    pitch = torch.randn(b, 1, t, generator=gen) if use_pitch else None
    energy = torch.randn(b, t, generator=gen) if use_energy else None
    return pitch, energy
