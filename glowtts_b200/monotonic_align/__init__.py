"""Drop-in for the reference's ``monotonic_align`` package (monotonic_align/__init__.py:6-21).

``maximum_path(value, mask)`` keeps the reference signature and result (0/1 alignment in the
input dtype on the input device) but never leaves the GPU: lengths come from the mask on the
device and the DP + backtrack run in one sm_100a kernel (csrc/mas.cu).
"""
from __future__ import annotations

import torch

from .. import ops


def maximum_path(value: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """value, mask: [b, t_x, t_y] (mask a rectangular 0/1 prefix mask) -> path [b, t_x, t_y]."""
    path, _, _ = maximum_path_full(value, mask)
    return path


def maximum_path_full(value: torch.Tensor, mask: torch.Tensor):
    """As maximum_path, plus durations int32 [b, t_x] (= path.sum(2), models.py:1085) and the
    per-frame token index int32 [b, t_y] (-1 in the padding) straight from the backtrack."""
    dtype = value.dtype
    with torch.no_grad():
        v = (value * mask).to(torch.float32)          # __init__.py:11,14
        t_xs, t_ys = ops.mas_lengths(mask.to(torch.float32))
        path, dur, idx = ops.mas_path(v, t_xs, t_ys)
    return path.to(dtype), dur, idx


def maximum_path_lengths(value: torch.Tensor, t_xs: torch.Tensor, t_ys: torch.Tensor):
    """core.pyx:40-45 entry (maximum_path_c) for callers that already hold the lengths:
    value fp32 [b,t_x,t_y] (already masked), t_xs / t_ys int32 [b] on the device."""
    with torch.no_grad():
        return ops.mas_path(value, t_xs, t_ys)


FUSED_MIN_BATCH = 96      # measured on B200 (profiles/ncu_mas_fused_r01.txt): below this the logp kernel's use of all SMs wins


def maximum_path_from_prior(z: torch.Tensor, x_m: torch.Tensor, x_logs, t_xs: torch.Tensor, t_ys: torch.Tensor,
                            return_logp: bool = False, fused=False):
    """models.py:1076-1083 as one call: the alignment of z [b,d,t_y] against the prior means x_m [b,d,t_x]
    (x_logs [b,d,t_x] or None for mean_only models), lengths int32 [b] on the device.
    -> (path, durations, index[, logp]).  Mean-only shapes the fused kernel covers (ops.mas_fused_supported) can run as
    ONE launch with logp computed on the tensor cores inside the MAS CTA -- OPT-IN (`fused=True`, or `fused="auto"`:
    when the batch has at least FUSED_MIN_BATCH utterances, i.e. when one CTA per utterance fills the GPU), because that
    kernel is not bit-exact with logp -> MAS (3xTF32 products, the per-frame z^2 term dropped from the DP): the default
    is the exact two-kernel path, so alignments never depend on the batch size."""
    with torch.no_grad():
        D, T_x, T_y = x_m.shape[1], x_m.shape[2], z.shape[2]
        can = ops.mas_fused_supported(D, T_x, T_y, x_logs is not None) and z.data_ptr() % 16 == 0
        if fused and not can:
            raise RuntimeError("glowtts_b200: this shape is not covered by the fused logp + MAS kernel")
        if can and (fused is True or (fused == "auto" and z.shape[0] >= FUSED_MIN_BATCH)):
            path, dur, idx, lp = ops.mas_fused_logp(z, x_m, t_xs, t_ys, return_logp)
        else:
            lp = ops.logp(x_m, x_logs, z)
            path, dur, idx = ops.mas_path(lp, t_xs, t_ys)
    return (path, dur, idx, lp) if return_logp else (path, dur, idx)
