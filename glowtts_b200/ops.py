"""torch.library ops over the C ABI (include/glowtts_b200.h).

Every op here is a thin shim: check dtype/device/contiguity, allocate outputs with torch,
pass raw device pointers and the current CUDA stream to libglowtts_b200.so.  No op has a CPU
implementation: calling one with CPU tensors (or without the built library) raises.
Ops are registered in the ``glowtts_b200`` namespace (``torch.ops.glowtts_b200.*``) with fake
(meta) implementations and autograd formulas, so they compose with autograd and DDP.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from . import _lib
from ._lib import BlockParams, DecoderDesc, GTK_MAX_LAYERS

NS = "glowtts_b200"


def _stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _chk(t: Tensor, name: str, dtype=torch.float32) -> Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"glowtts_b200: {name} must be a CUDA tensor (there is no CPU path)")
    if t.dtype != dtype:
        raise RuntimeError(f"glowtts_b200: {name} must be {dtype}, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


def _opt(t: Optional[Tensor], name: str) -> Optional[Tensor]:
    return None if t is None else _chk(t, name)


# =============================================================================== MAS
@torch.library.custom_op(f"{NS}::mas_lengths", mutates_args=(), device_types="cuda")
def mas_lengths(mask: Tensor) -> Tuple[Tensor, Tensor]:
    """monotonic_align/__init__.py:18-19 on the device: (t_x, t_y) int32 [B] from a prefix mask."""
    mask = _chk(mask, "mask")
    B, T_x, T_y = mask.shape
    t_xs = torch.empty(B, dtype=torch.int32, device=mask.device)
    t_ys = torch.empty(B, dtype=torch.int32, device=mask.device)
    _lib.call("gtk_mas_lengths_from_mask", _p(mask), _p(t_xs), _p(t_ys), B, T_x, T_y, _stream())
    return t_xs, t_ys


@mas_lengths.register_fake
def _(mask):
    B = mask.shape[0]
    return mask.new_empty(B, dtype=torch.int32), mask.new_empty(B, dtype=torch.int32)


@torch.library.custom_op(f"{NS}::mas_path", mutates_args=(), device_types="cuda")
def mas_path(value: Tensor, t_xs: Tensor, t_ys: Tensor, force_generic: int = 0) -> Tuple[Tensor, Tensor, Tensor]:
    """core.pyx:9-45 on the device.  -> (path f32 [B,T_x,T_y], durations i32 [B,T_x], index i32 [B,T_y]).
    force_generic: 0 = fastest applicable kernel (TMA ring), 1 = generic kernel, 2 = helper-warp fast path."""
    value = _chk(value, "value")
    t_xs = _chk(t_xs, "t_xs", torch.int32)
    t_ys = _chk(t_ys, "t_ys", torch.int32)
    B, T_x, T_y = value.shape
    path = torch.empty_like(value)
    dur = torch.empty(B, T_x, dtype=torch.int32, device=value.device)
    idx = torch.empty(B, T_y, dtype=torch.int32, device=value.device)
    lib = _lib.load()
    ws_bytes = 0 if not force_generic else B * T_y * ((T_x + 31) // 32) * 4
    ws_bytes = max(ws_bytes, lib.gtk_mas_workspace_bytes(B, T_x, T_y))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=value.device) if ws_bytes else None
    _lib.call("gtk_mas_forward", _p(value), _p(t_xs), _p(t_ys), _p(path), _p(dur), _p(idx), B, T_x, T_y,
              _p(ws), ws_bytes, int(force_generic), _stream())
    return path, dur, idx


@mas_path.register_fake
def _(value, t_xs, t_ys, force_generic=0):
    B, T_x, T_y = value.shape
    return (torch.empty_like(value), value.new_empty((B, T_x), dtype=torch.int32),
            value.new_empty((B, T_y), dtype=torch.int32))


@torch.library.custom_op(f"{NS}::mas_index", mutates_args=(), device_types="cuda")
def mas_index(value: Tensor, t_xs: Tensor, t_ys: Tensor) -> Tuple[Tensor, Tensor]:
    """core.pyx:9-45 without the [B,T_x,T_y] path plane (gtk_mas_forward with path = NULL): -> (durations i32 [B,T_x],
    index i32 [B,T_y]).  What the training step needs (models.py:1085, 1118-1119 as a gather): the float path is
    never written, so the kernel's HBM traffic is the 4 B/cell read of `value`."""
    value = _chk(value, "value")
    t_xs = _chk(t_xs, "t_xs", torch.int32)
    t_ys = _chk(t_ys, "t_ys", torch.int32)
    B, T_x, T_y = value.shape
    dur = torch.empty(B, T_x, dtype=torch.int32, device=value.device)
    idx = torch.empty(B, T_y, dtype=torch.int32, device=value.device)
    ws_bytes = _lib.load().gtk_mas_workspace_bytes(B, T_x, T_y)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=value.device) if ws_bytes else None
    _lib.call("gtk_mas_forward", _p(value), _p(t_xs), _p(t_ys), None, _p(dur), _p(idx), B, T_x, T_y, _p(ws), ws_bytes, 0, _stream())
    return dur, idx


@mas_index.register_fake
def _(value, t_xs, t_ys):
    B, T_x, T_y = value.shape
    return value.new_empty((B, T_x), dtype=torch.int32), value.new_empty((B, T_y), dtype=torch.int32)


def mas_fused_supported(D: int, T_x: int, T_y: int, has_x_logs: bool) -> bool:
    """Whether gtk_mas_fused_logp covers this shape (mean_only, D == 80, T_x <= 256, T_y % 4 == 0)."""
    return bool(_lib.load().gtk_mas_fused_supported(int(D), int(T_x), int(T_y), int(has_x_logs)))


@torch.library.custom_op(f"{NS}::mas_fused_logp", mutates_args=(), device_types="cuda")
def mas_fused_logp(z: Tensor, x_m: Tensor, t_xs: Tensor, t_ys: Tensor, want_logp: bool = False) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """models.py:1076-1083 in one kernel (mean_only): logp is built tile by tile inside the MAS CTA and never stored.
    z [B,D,T_y], x_m [B,D,T_x] -> (path, durations, index, logp) -- logp is [B,T_x,T_y] (zero outside the lengths) when
    want_logp, else an empty tensor.  Raises for shapes the fused kernel does not cover (see mas_fused_supported)."""
    z = _chk(z, "z")
    x_m = _chk(x_m, "x_m")
    t_xs = _chk(t_xs, "t_xs", torch.int32)
    t_ys = _chk(t_ys, "t_ys", torch.int32)
    B, D, T_x = x_m.shape
    T_y = z.shape[2]
    path = torch.empty(B, T_x, T_y, dtype=torch.float32, device=z.device)
    dur = torch.empty(B, T_x, dtype=torch.int32, device=z.device)
    idx = torch.empty(B, T_y, dtype=torch.int32, device=z.device)
    lp = torch.zeros(B, T_x, T_y, dtype=torch.float32, device=z.device) if want_logp else z.new_empty(0)
    _lib.call("gtk_mas_fused_logp", _p(z), _p(x_m), None, _p(t_xs), _p(t_ys), _p(path), _p(dur), _p(idx),
              _p(lp) if want_logp else None, B, D, T_x, T_y, _stream())
    return path, dur, idx, lp


@mas_fused_logp.register_fake
def _(z, x_m, t_xs, t_ys, want_logp=False):
    B, D, T_x = x_m.shape
    T_y = z.shape[2]
    return (z.new_empty((B, T_x, T_y)), z.new_empty((B, T_x), dtype=torch.int32), z.new_empty((B, T_y), dtype=torch.int32),
            z.new_empty((B, T_x, T_y)) if want_logp else z.new_empty(0))


# =============================================================================== clip_grad_value_
@torch.library.custom_op(f"{NS}::clip_grad_value", mutates_args=("grad", "acc"), device_types="cuda")
def clip_grad_value(grad: Tensor, clip_value: float, do_clip: bool, norm_type: float, acc: Tensor) -> None:
    """commons.py:320-336 over one contiguous gradient range: acc += sum |g|^p, clamp in place if do_clip."""
    if not (grad.is_cuda and grad.dtype == torch.float32 and grad.is_contiguous()):
        raise RuntimeError("glowtts_b200: clip_grad_value needs a contiguous CUDA fp32 gradient")
    acc = _chk(acc, "acc")
    _lib.call("gtk_clip_grad_value", _p(grad), grad.numel(), float(clip_value), int(do_clip), float(norm_type), _p(acc),
              _stream())


# =============================================================================== logp
@torch.library.custom_op(f"{NS}::logp", mutates_args=(), device_types="cuda")
def logp(x_m: Tensor, x_logs: Optional[Tensor], z: Tensor) -> Tensor:
    """models.py:1076-1082.  x_m, x_logs [B,D,T_x]; z [B,D,T_y] -> [B,T_x,T_y].  x_logs None == 0."""
    x_m = _chk(x_m, "x_m")
    x_logs = _opt(x_logs, "x_logs")
    z = _chk(z, "z")
    B, D, T_x = x_m.shape
    T_y = z.shape[2]
    out = torch.empty(B, T_x, T_y, dtype=torch.float32, device=z.device)
    _lib.call("gtk_logp", _p(x_m), _p(x_logs), _p(z), _p(out), B, D, T_x, T_y, _stream())
    return out


@logp.register_fake
def _(x_m, x_logs, z):
    return z.new_empty((x_m.shape[0], x_m.shape[2], z.shape[2]))


# =============================================================================== glue
@torch.library.custom_op(f"{NS}::squeeze", mutates_args=(), device_types="cuda")
def squeeze_op(x: Tensor, mask: Optional[Tensor], n_sqz: int) -> Tuple[Tensor, Tensor]:
    """commons.py:339-351.  x [B,C,T], mask [B,T] -> (x_sqz [B,C*n,T//n], mask_sqz [B,T//n])."""
    x = _chk(x, "x")
    mask = _opt(mask, "mask")
    B, C, T = x.shape
    Tq = T // n_sqz
    out = torch.empty(B, C * n_sqz, Tq, dtype=x.dtype, device=x.device)
    mo = torch.empty(B, Tq, dtype=x.dtype, device=x.device)
    _lib.call("gtk_squeeze", _p(x), _p(mask), _p(out), _p(mo), B, C, T, n_sqz, _stream())
    return out, mo


@squeeze_op.register_fake
def _(x, mask, n_sqz):
    B, C, T = x.shape
    return x.new_empty((B, C * n_sqz, T // n_sqz)), x.new_empty((B, T // n_sqz))


@torch.library.custom_op(f"{NS}::unsqueeze", mutates_args=(), device_types="cuda")
def unsqueeze_op(x: Tensor, mask: Optional[Tensor], n_sqz: int) -> Tuple[Tensor, Tensor]:
    """commons.py:354-364.  x [B,C,T], mask [B,T] -> (x_unsqz [B,C//n,T*n], mask [B,T*n])."""
    x = _chk(x, "x")
    mask = _opt(mask, "mask")
    B, C, T = x.shape
    out = torch.empty(B, C // n_sqz, T * n_sqz, dtype=x.dtype, device=x.device)
    mo = torch.empty(B, T * n_sqz, dtype=x.dtype, device=x.device)
    _lib.call("gtk_unsqueeze", _p(x), _p(mask), _p(out), _p(mo), B, C, T, n_sqz, _stream())
    return out, mo


@unsqueeze_op.register_fake
def _(x, mask, n_sqz):
    B, C, T = x.shape
    return x.new_empty((B, C // n_sqz, T * n_sqz)), x.new_empty((B, T * n_sqz))


def _squeeze_bwd(ctx, g_out, g_mask):
    # squeeze and unsqueeze are inverse index maps; the output mask is applied in both.
    dx, _ = unsqueeze_op(g_out, ctx.mask_out, ctx.n_sqz)
    if ctx.T != dx.shape[2]:
        dx = torch.nn.functional.pad(dx, (0, ctx.T - dx.shape[2]))
    return dx, None, None


def _squeeze_setup(ctx, inputs, output):
    ctx.n_sqz = inputs[2]
    ctx.T = inputs[0].shape[2]
    ctx.mask_out = output[1] if inputs[1] is not None else None


squeeze_op.register_autograd(_squeeze_bwd, setup_context=_squeeze_setup)


def _unsqueeze_bwd(ctx, g_out, g_mask):
    dx, _ = squeeze_op(g_out, ctx.mask_out, ctx.n_sqz)
    return dx, None, None


def _unsqueeze_setup(ctx, inputs, output):
    ctx.n_sqz = inputs[2]
    ctx.mask_out = output[1] if inputs[1] is not None else None


unsqueeze_op.register_autograd(_unsqueeze_bwd, setup_context=_unsqueeze_setup)


# ---- gate (commons.py:61-68)
def _cond_mode(a: Tensor, b: Optional[Tensor]) -> int:
    if b is None:
        return 0
    if b.dim() == 3 and b.shape[2] == a.shape[2] and a.shape[2] != 1:
        return 2
    if b.shape[-1] == 1 or b.dim() == 2:
        return 1
    return 2


@torch.library.custom_op(f"{NS}::gate_fwd", mutates_args=(), device_types="cuda")
def gate_fwd(a: Tensor, b: Optional[Tensor]) -> Tensor:
    a = _chk(a, "input_a")
    b = _opt(b, "input_b")
    B, H2, T = a.shape
    out = torch.empty(B, H2 // 2, T, dtype=a.dtype, device=a.device)
    _lib.call("gtk_gate_fwd", _p(a), _p(b), _cond_mode(a, b), _p(out), B, H2 // 2, T, _stream())
    return out


@gate_fwd.register_fake
def _(a, b):
    return a.new_empty((a.shape[0], a.shape[1] // 2, a.shape[2]))


@torch.library.custom_op(f"{NS}::gate_bwd", mutates_args=(), device_types="cuda")
def gate_bwd(a: Tensor, b: Optional[Tensor], dout: Tensor) -> Tensor:
    a = _chk(a, "input_a")
    b = _opt(b, "input_b")
    dout = _chk(dout, "dout")
    B, H2, T = a.shape
    da = torch.empty_like(a)
    _lib.call("gtk_gate_bwd", _p(a), _p(b), _cond_mode(a, b), _p(dout), _p(da), B, H2 // 2, T, _stream())
    return da


@gate_bwd.register_fake
def _(a, b, dout):
    return torch.empty_like(a)


@torch.library.custom_op(f"{NS}::row_sum", mutates_args=(), device_types="cuda")
def row_sum(x: Tensor) -> Tensor:
    """Sum over the last dimension of a contiguous tensor (bias / broadcast-conditioner gradients)."""
    x = _chk(x, "x")
    T = x.shape[-1]
    R = x.numel() // max(T, 1)
    out = torch.empty(x.shape[:-1], dtype=x.dtype, device=x.device)
    _lib.call("gtk_row_sum", _p(x), _p(out), R, T, 0, _stream())
    return out


@row_sum.register_fake
def _(x):
    return x.new_empty(x.shape[:-1])


def _gate_setup(ctx, inputs, output):
    a, b = inputs
    ctx.save_for_backward(a, b)
    ctx.b_shape = None if b is None else tuple(b.shape)


def _gate_bwd(ctx, dout):
    a, b = ctx.saved_tensors
    da = gate_bwd(a, b, dout.contiguous())
    db = None
    if b is not None and ctx.needs_input_grad[1]:
        db = da if _cond_mode(a, b) == 2 else row_sum(da).reshape(ctx.b_shape)
    return da, db


gate_fwd.register_autograd(_gate_bwd, setup_context=_gate_setup)


# =============================================================================== ActNorm / InvConvNear
@torch.library.custom_op(f"{NS}::mask_len", mutates_args=(), device_types="cuda")
def mask_len(mask: Tensor) -> Tensor:
    """x_len[b] = sum_t mask[b,t] (modules.py:587,642).  mask [B,T]."""
    mask = _chk(mask, "mask")
    B, T = mask.shape
    out = torch.empty(B, dtype=mask.dtype, device=mask.device)
    _lib.call("gtk_mask_len", _p(mask), _p(out), B, T, _stream())
    return out


@mask_len.register_fake
def _(mask):
    return mask.new_empty((mask.shape[0],))


@torch.library.custom_op(f"{NS}::invconv_prep", mutates_args=(), device_types="cuda")
def invconv_prep(weight: Tensor) -> Tuple[Tensor, Tensor]:
    """(W^-1, logdet W) of the n_split x n_split InvConvNear weight (modules.py:651,658,668), on device."""
    weight = _chk(weight, "weight")
    ns = weight.shape[0]
    winv = torch.empty_like(weight)
    ld = torch.empty((), dtype=weight.dtype, device=weight.device)
    _lib.call("gtk_invconv_prep", _p(weight), ns, _p(winv), _p(ld), _stream())
    return winv, ld


@invconv_prep.register_fake
def _(weight):
    return torch.empty_like(weight), weight.new_empty(())


@torch.library.custom_op(f"{NS}::flow_mix", mutates_args=(), device_types="cuda")
def flow_mix(x: Tensor, mask: Optional[Tensor], logs: Optional[Tensor], bias: Optional[Tensor], M: Optional[Tensor],
             n_split: int, reverse: bool) -> Tensor:
    """Fused ActNorm (+) InvConvNear (gtk_flow_mix).  x [B,C,T], mask [B,T], logs/bias [C], M [ns,ns]."""
    x = _chk(x, "x")
    mask, logs, bias, M = _opt(mask, "mask"), _opt(logs, "logs"), _opt(bias, "bias"), _opt(M, "M")
    B, C, T = x.shape
    z = torch.empty_like(x)
    _lib.call("gtk_flow_mix", _p(x), _p(mask), _p(logs), _p(bias), _p(M), n_split, _p(z), B, C, T, int(reverse), _stream())
    return z


@flow_mix.register_fake
def _(x, mask, logs, bias, M, n_split, reverse):
    return torch.empty_like(x)


@torch.library.custom_op(f"{NS}::flow_mix_bwd", mutates_args=(), device_types="cuda")
def flow_mix_bwd(x: Tensor, mask: Optional[Tensor], logs: Optional[Tensor], bias: Optional[Tensor], M: Optional[Tensor],
                 n_split: int, dz: Tensor) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """-> (dx, dlogs [C], dbias [C], dM [ns,ns]) of the forward mix."""
    x, dz = _chk(x, "x"), _chk(dz, "dz")
    mask, logs, bias, M = _opt(mask, "mask"), _opt(logs, "logs"), _opt(bias, "bias"), _opt(M, "M")
    B, C, T = x.shape
    dx = torch.empty_like(x)
    dlogs = torch.zeros(C, dtype=x.dtype, device=x.device)
    dbias = torch.zeros(C, dtype=x.dtype, device=x.device)
    dM = torch.zeros(n_split, n_split, dtype=x.dtype, device=x.device)
    _lib.call("gtk_flow_mix_bwd", _p(x), _p(mask), _p(logs), _p(bias), _p(M), n_split, _p(dz), _p(dx), _p(dlogs), _p(dbias),
              _p(dM), B, C, T, _stream())
    return dx, dlogs, dbias, dM


@flow_mix_bwd.register_fake
def _(x, mask, logs, bias, M, n_split, dz):
    C = x.shape[1]
    return torch.empty_like(x), x.new_empty((C,)), x.new_empty((C,)), x.new_empty((n_split, n_split))


def _mix_setup(ctx, inputs, output):
    x, mask, logs, bias, M, n_split, reverse = inputs
    ctx.save_for_backward(x, mask, logs, bias, M)
    ctx.n_split, ctx.reverse = n_split, reverse


def _mix_bwd(ctx, dz):
    x, mask, logs, bias, M = ctx.saved_tensors
    if ctx.reverse:
        raise RuntimeError("glowtts_b200: the reverse flow is inference-only (no backward), as in the reference's use")
    dx, dlogs, dbias, dM = flow_mix_bwd(x, mask, logs, bias, M, ctx.n_split, dz.contiguous())
    return (dx, None, dlogs if logs is not None else None, dbias if bias is not None else None,
            dM if M is not None else None, None, None)


flow_mix.register_autograd(_mix_bwd, setup_context=_mix_setup)


@torch.library.custom_op(f"{NS}::flow_logdet_const", mutates_args=(), device_types="cuda")
def flow_logdet_const(logs: Optional[Tensor], logdet_w: Optional[Tensor], group_count: float, x_len: Tensor) -> Tensor:
    """logdet[b] = (sum_c logs_c + logdet_w * group_count) * x_len[b]  (modules.py:597, 658)."""
    logs, logdet_w = _opt(logs, "logs"), _opt(logdet_w, "logdet_w")
    x_len = _chk(x_len, "x_len")
    B = x_len.shape[0]
    out = torch.empty(B, dtype=x_len.dtype, device=x_len.device)
    C = 0 if logs is None else logs.numel()
    _lib.call("gtk_flow_logdet_const", _p(logs), C, _p(logdet_w), ctypes.c_float(group_count), _p(x_len), _p(out), B, 0,
              _stream())
    return out


@flow_logdet_const.register_fake
def _(logs, logdet_w, group_count, x_len):
    return torch.empty_like(x_len)


@torch.library.custom_op(f"{NS}::flow_logdet_const_bwd", mutates_args=(), device_types="cuda")
def flow_logdet_const_bwd(dlogdet: Tensor, x_len: Tensor, C: int, winv: Optional[Tensor], group_count: float) -> Tuple[Tensor, Tensor]:
    """-> (dlogs [C] (all equal), dW [ns,ns] = W^-T * group_count * sum_b dlogdet_b x_len_b)."""
    dlogdet, x_len = _chk(dlogdet, "dlogdet"), _chk(x_len, "x_len")
    winv = _opt(winv, "winv")
    B = x_len.shape[0]
    dlogs = torch.zeros(max(C, 1), dtype=x_len.dtype, device=x_len.device)
    ns = 1 if winv is None else winv.shape[0]
    dW = torch.zeros(ns, ns, dtype=x_len.dtype, device=x_len.device)
    _lib.call("gtk_flow_logdet_const_bwd", _p(dlogdet), _p(x_len), B, _p(dlogs) if C else None, C, _p(winv), ns,
              ctypes.c_float(group_count), _p(dW) if winv is not None else None, _stream())
    return dlogs, dW


@flow_logdet_const_bwd.register_fake
def _(dlogdet, x_len, C, winv, group_count):
    ns = 1 if winv is None else winv.shape[0]
    return x_len.new_empty((max(C, 1),)), x_len.new_empty((ns, ns))


@torch.library.custom_op(f"{NS}::actnorm_ddi", mutates_args=(), device_types="cuda")
def actnorm_ddi_stats(x: Tensor, mask: Optional[Tensor]) -> Tensor:
    """ActNorm.initialize statistics (modules.py:609-611): [3,C] = (sum mask, sum x, sum x^2)."""
    x = _chk(x, "x")
    mask = _opt(mask, "mask")
    B, C, T = x.shape
    stats = torch.zeros(3, C, dtype=x.dtype, device=x.device)
    _lib.call("gtk_actnorm_ddi_stats", _p(x), _p(mask), _p(stats), B, C, T, _stream())
    return stats


@actnorm_ddi_stats.register_fake
def _(x, mask):
    return x.new_empty((3, x.shape[1]))


def actnorm_ddi_finalize(stats: Tensor) -> Tuple[Tensor, Tensor]:
    """modules.py:612-619: stats -> (logs [C], bias [C])."""
    stats = _chk(stats, "stats")
    C = stats.shape[1]
    logs = torch.empty(C, dtype=stats.dtype, device=stats.device)
    bias = torch.empty(C, dtype=stats.dtype, device=stats.device)
    _lib.call("gtk_actnorm_ddi_finalize", _p(stats), _p(logs), _p(bias), C, _stream())
    return logs, bias


# =============================================================================== loss / alignment consumers
@torch.library.custom_op(f"{NS}::mle_loss_fwd", mutates_args=(), device_types="cuda")
def mle_loss_fwd(z: Tensor, m: Optional[Tensor], logs: Optional[Tensor], logdet: Tensor, mask: Tensor) -> Tuple[Tensor, Tensor]:
    """commons.py:28-33 -> (loss scalar, acc4 scratch kept for the backward).  mask [B,T]."""
    z, logdet, mask = _chk(z, "z"), _chk(logdet, "logdet"), _chk(mask, "mask")
    m, logs = _opt(m, "m"), _opt(logs, "logs")
    B, C, T = z.shape
    acc = torch.empty(4, dtype=z.dtype, device=z.device)
    loss = torch.empty((), dtype=z.dtype, device=z.device)
    _lib.call("gtk_mle_loss_fwd", _p(z), _p(m), _p(logs), _p(logdet), _p(mask), _p(acc), _p(loss), B, C, T, _stream())
    return loss, acc


@mle_loss_fwd.register_fake
def _(z, m, logs, logdet, mask):
    return z.new_empty(()), z.new_empty((4,))


@torch.library.custom_op(f"{NS}::mle_loss_bwd", mutates_args=(), device_types="cuda")
def mle_loss_bwd(z: Tensor, m: Optional[Tensor], logs: Optional[Tensor], acc: Tensor, gloss: Tensor,
                 n_batch: int) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    z, acc, gloss = _chk(z, "z"), _chk(acc, "acc4"), _chk(gloss, "gloss")
    m, logs = _opt(m, "m"), _opt(logs, "logs")
    B, C, T = z.shape
    dz = torch.empty_like(z)
    dm = torch.empty_like(z) if m is not None else z.new_empty(0)
    dlogs = torch.empty_like(z) if logs is not None else z.new_empty(0)
    dld = torch.empty(n_batch, dtype=z.dtype, device=z.device)
    _lib.call("gtk_mle_loss_bwd", _p(z), _p(m), _p(logs), _p(acc), _p(gloss), _p(dz), _p(dm) if m is not None else None,
              _p(dlogs) if logs is not None else None, _p(dld), B, C, T, _stream())
    return dz, dm, dlogs, dld


@mle_loss_bwd.register_fake
def _(z, m, logs, acc, gloss, n_batch):
    return (torch.empty_like(z), torch.empty_like(z) if m is not None else z.new_empty(0),
            torch.empty_like(z) if logs is not None else z.new_empty(0), z.new_empty((n_batch,)))


def _mle_setup(ctx, inputs, output):
    z, m, logs, logdet, mask = inputs
    ctx.save_for_backward(z, m, logs, output[1])
    ctx.n_batch = logdet.shape[0]


def _mle_bwd(ctx, gloss, gacc):
    z, m, logs, acc = ctx.saved_tensors
    dz, dm, dlogs, dld = mle_loss_bwd(z, m, logs, acc, gloss.contiguous(), ctx.n_batch)
    return dz, (dm if m is not None else None), (dlogs if logs is not None else None), dld, None


mle_loss_fwd.register_autograd(_mle_bwd, setup_context=_mle_setup)


@torch.library.custom_op(f"{NS}::gather_tokens", mutates_args=(), device_types="cuda")
def gather_tokens(src: Tensor, idx: Tensor, T_y: int) -> Tensor:
    """models.py:1118-1119 as a gather: out[b,c,y] = src[b,c,idx[b,y]] (0 where idx < 0).  src [B,C,T_x]."""
    src, idx = _chk(src, "src"), _chk(idx, "idx", torch.int32)
    B, C, T_x = src.shape
    out = torch.empty(B, C, T_y, dtype=src.dtype, device=src.device)
    _lib.call("gtk_gather_tokens", _p(src), _p(idx), _p(out), B, C, T_x, T_y, _stream())
    return out


@gather_tokens.register_fake
def _(src, idx, T_y):
    return src.new_empty((src.shape[0], src.shape[1], T_y))


def _gather_setup(ctx, inputs, output):
    ctx.save_for_backward(inputs[1])
    ctx.T_x = inputs[0].shape[2]


def _gather_bwd(ctx, dout):
    (idx,) = ctx.saved_tensors
    # scatter-add of a monotone index map: small ([B,C,T_x]); library op is fine off the hot path
    B, C, T_y = dout.shape
    valid = (idx >= 0)
    ix = idx.clamp(min=0).long().unsqueeze(1).expand(B, C, T_y)
    dsrc = torch.zeros(B, C, ctx.T_x, dtype=dout.dtype, device=dout.device)
    dsrc.scatter_add_(2, ix, dout * valid.unsqueeze(1))
    return dsrc, None, None


gather_tokens.register_autograd(_gather_bwd, setup_context=_gather_setup)


@torch.library.custom_op(f"{NS}::generate_path", mutates_args=(), device_types="cuda")
def generate_path(duration: Tensor, mask: Tensor) -> Tensor:
    """commons.py:127-143.  duration [B,T_x] float, mask [B,T_x,T_y] -> path [B,T_x,T_y]."""
    duration, mask = _chk(duration, "duration"), _chk(mask, "mask")
    B, T_x, T_y = mask.shape
    path = torch.empty_like(mask)
    _lib.call("gtk_generate_path", _p(duration), _p(mask), _p(path), B, T_x, T_y, _stream())
    return path


@generate_path.register_fake
def _(duration, mask):
    return torch.empty_like(mask)


# =============================================================================== whole decoder
# cfg = [in_channels, n_sqz, hidden, kernel_size, dilation_rate, n_layers, n_blocks, n_split, sigmoid_scale,
#        precision, wn_folded, p_dropout (ppm), host seed, prosody]
# params: per block, in this order (g entries absent when wn_folded):
#   an_logs, an_bias, ic_weight, start_g, start_v, start_b, in_g*L, in_v*L, in_b*L, rs_g*L, rs_v*L, rs_b*L, end_w, end_b
# and, when cfg[13] (prosody: the fork's wn_energy / wn_pitch WaveNets, attentions.py:114-117, 155-160) is set, after end_b:
#   for wn_energy, then wn_pitch:  in_g*L, in_v*L, in_b*L, rs_g*L, rs_v*L, rs_b*L, cond_layer1 g, v, b
CFG_LEN = 11


def params_per_wnp(n_layers: int, wn_folded: bool) -> int:
    return (4 * n_layers + 2) if wn_folded else (6 * n_layers + 3)


def params_per_block(n_layers: int, wn_folded: bool, prosody: bool = False) -> int:
    base = (7 + 4 * n_layers) if wn_folded else (8 + 6 * n_layers)
    return base + (2 * params_per_wnp(n_layers, wn_folded) if prosody else 0)


def _prosody(cfg: Sequence[int]) -> bool:
    return len(cfg) > 13 and bool(cfg[13])


def _track(t: Optional[Tensor], name: str, B: int, T: int) -> Optional[Tensor]:
    """A pitch / energy track as the library reads it: fp32 [B, T] (a [B,1,T'] track with T' >= T is cut, models.py:1059)."""
    if t is None:
        return None
    t = t.reshape(t.shape[0], t.shape[-1])
    if t.shape[0] != B or t.shape[1] < T:
        raise RuntimeError(f"glowtts_b200: {name} must be [B, T] with B={B}, T>={T}; got {tuple(t.shape)}")
    if t.shape[1] != T:
        t = t[:, :T]
    return _chk(t.contiguous(), name)


def _desc(cfg: Sequence[int], B: int, T: int, seed: Optional[Tensor] = None, energy: Optional[Tensor] = None,
          pitch: Optional[Tensor] = None, d_energy: Optional[Tensor] = None, d_pitch: Optional[Tensor] = None) -> DecoderDesc:
    d = DecoderDesc()
    if (energy is not None or pitch is not None) and not _prosody(cfg):
        raise RuntimeError("glowtts_b200: pitch / energy tracks need cfg[13] = 1 and the wn_energy / wn_pitch parameters")
    d.energy, d.pitch, d.d_energy, d.d_pitch = _p(energy), _p(pitch), _p(d_energy), _p(d_pitch)
    d.B, d.T = B, T
    (d.in_channels, d.n_sqz, d.hidden, d.kernel_size, d.dilation_rate, d.n_layers, d.n_blocks, d.n_split,
     d.sigmoid_scale, d.precision) = [int(v) for v in cfg[:10]]
    if len(cfg) > 12:                      # cfg[11] = dropout probability in ppm, cfg[12] = per-call seed (host value)
        d.p_dropout, d.seed = int(cfg[11]) * 1e-6, int(cfg[12]) & 0xFFFFFFFFFFFFFFFF
    if seed is not None:                   # device-resident per-call seed (CUDA-graph friendly); overrides cfg[12]
        if seed.dtype != torch.int64 or seed.numel() != 1 or not seed.is_cuda:
            raise RuntimeError("glowtts_b200: dropout seed must be a one-element int64 CUDA tensor")
        d.seed_dev = seed.data_ptr()
    if d.n_layers > GTK_MAX_LAYERS:
        raise RuntimeError(f"glowtts_b200: n_layers {d.n_layers} > {GTK_MAX_LAYERS}")
    return d


def draw_seed() -> int:
    """A fresh dropout seed from torch's CPU generator (reproducible under torch.manual_seed, no device sync)."""
    return int(torch.randint(0, 2 ** 62, (1,)).item())


def dropout_scale(seed: int, block: int, layer: int, p: float, B: int, T: int, C2: int, device) -> Tensor:
    """The dropout keep-scale (0 or 1/(1-p)) the decoder kernels apply to the in_layer output of (block, layer)
    (modules.py:153) for per-call seed `seed`, as a [B, C2, T] tensor."""
    lib = _lib.load()
    lseed = int(lib.gtk_decoder_layer_seed(int(seed) & 0xFFFFFFFFFFFFFFFF, int(block), int(layer)))
    out = torch.empty(B, T, C2, dtype=torch.float32, device=device)
    _lib.call("gtk_dropout_scale", lseed, float(p), out.numel(), _p(out), _stream())
    return out.transpose(1, 2)


def _blocks(cfg: Sequence[int], tensors: Sequence[Optional[Tensor]], cond: Optional[Tensor]):
    """Fill an array of gtk_block_params (or gtk_block_grads) from the flat tensor list."""
    L, NB, folded = int(cfg[5]), int(cfg[6]), bool(cfg[10])
    per = params_per_block(L, folded, _prosody(cfg))
    if len(tensors) != per * NB:
        raise RuntimeError(f"glowtts_b200: expected {per * NB} decoder tensors, got {len(tensors)}")
    arr = (BlockParams * NB)()
    ptr = lambda t: None if t is None else (t if isinstance(t, int) else t.data_ptr())
    for b in range(NB):
        it = iter(tensors[b * per:(b + 1) * per])
        P = arr[b]
        P.an_logs, P.an_bias, P.ic_weight = ptr(next(it)), ptr(next(it)), ptr(next(it))
        P.start_g = None if folded else ptr(next(it))
        P.start_v, P.start_b = ptr(next(it)), ptr(next(it))
        for name in ("in_g", "in_v", "in_b", "rs_g", "rs_v", "rs_b"):
            fld = getattr(P, name)
            for i in range(L):
                fld[i] = None if (folded and name.endswith("_g")) else ptr(next(it))
        P.end_w, P.end_b = ptr(next(it)), ptr(next(it))
        if _prosody(cfg):
            for W in (P.wn_energy, P.wn_pitch):
                for name in ("in_g", "in_v", "in_b", "rs_g", "rs_v", "rs_b"):
                    fld = getattr(W, name)
                    for i in range(L):
                        fld[i] = None if (folded and name.endswith("_g")) else ptr(next(it))
                W.c1_g = None if folded else ptr(next(it))
                W.c1_v, W.c1_b = ptr(next(it)), ptr(next(it))
        P.cond = None if cond is None else cond[b].data_ptr()
    return arr


def grad_layout(params: Sequence[Tensor]):
    """Offsets (16-byte aligned), sizes and total length of the flat gradient buffer of decoder_backward."""
    offs, sizes, tot = [], [], 0
    for p in params:
        n = p.numel()
        offs.append(tot)
        sizes.append(n)
        tot += (n + 3) // 4 * 4
    return offs, sizes, tot


def _fit_mask(mask: Tensor, T: int) -> Tensor:
    """The library reads mask rows with the data tensor's row length; a longer mask (the reference squeezes a
    [B,1,T] mask against a tensor already cut to a multiple of n_sqz, commons.py:342-348) is cut to match."""
    if mask.shape[-1] < T:
        raise RuntimeError(f"glowtts_b200: mask has {mask.shape[-1]} frames, data has {T}")
    return mask if mask.shape[-1] == T else mask[..., :T]


def _chk_list(ts: Sequence[Tensor], name: str) -> List[Tensor]:
    return [_chk(t, f"{name}[{i}]") for i, t in enumerate(ts)]


@torch.library.custom_op(f"{NS}::decoder_forward", mutates_args=(), device_types="cuda")
def decoder_forward(x: Tensor, mask: Tensor, params: List[Tensor], cond: Optional[Tensor], cfg: List[int],
                    train: bool, seed: Optional[Tensor] = None, energy: Optional[Tensor] = None,
                    pitch: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor]:
    """FlowSpecDecoder.forward(reverse=False) (models.py:765-785).  x [B,in,T], mask [B,T], cond [NB,B,2H*L] or None,
    energy / pitch [B,T] or None (the fork's per-frame tracks, bf16 path) -> (z [B,in,(T//n)*n], logdet [B], workspace
    kept for the backward)."""
    x = _chk(x, "x")
    B, Cin, T = x.shape
    mask = _chk(_fit_mask(mask, T), "mask")
    params = _chk_list(params, "params")
    cond = _opt(cond, "cond")
    d = _desc(cfg, B, T, seed, _track(energy, "energy", B, T), _track(pitch, "pitch", B, T))
    lib = _lib.load()
    ws_bytes = lib.gtk_decoder_workspace_bytes(ctypes.byref(d), int(train))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    Tz = (T // d.n_sqz) * d.n_sqz
    z = torch.empty(B, Cin, Tz, dtype=x.dtype, device=x.device)
    logdet = torch.empty(B, dtype=x.dtype, device=x.device)
    blocks = _blocks(cfg, params, cond)
    _lib.call("gtk_decoder_forward", ctypes.byref(d), blocks, _p(x), _p(mask), _p(z), _p(logdet), _p(ws), ws_bytes,
              int(train), _stream())
    if not train:
        ws = ws.new_empty(0)
    return z, logdet, ws


@decoder_forward.register_fake
def _(x, mask, params, cond, cfg, train, seed=None, energy=None, pitch=None):
    B, Cin, T = x.shape
    n = cfg[1]
    return x.new_empty((B, Cin, (T // n) * n)), x.new_empty((B,)), x.new_empty((0,), dtype=torch.uint8)


@torch.library.custom_op(f"{NS}::decoder_reverse", mutates_args=(), device_types="cuda")
def decoder_reverse(z: Tensor, mask: Tensor, params: List[Tensor], cond: Optional[Tensor], cfg: List[int],
                    seed: Optional[Tensor] = None, energy: Optional[Tensor] = None, pitch: Optional[Tensor] = None) -> Tensor:
    """FlowSpecDecoder.forward(reverse=True) (models.py:765-785)."""
    z = _chk(z, "z")
    B, Cin, T = z.shape
    mask = _chk(_fit_mask(mask, T), "mask")
    params = _chk_list(params, "params")
    cond = _opt(cond, "cond")
    d = _desc(cfg, B, T, seed, _track(energy, "energy", B, T), _track(pitch, "pitch", B, T))
    lib = _lib.load()
    ws_bytes = lib.gtk_decoder_workspace_bytes(ctypes.byref(d), 0)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=z.device)
    Tz = (T // d.n_sqz) * d.n_sqz
    x = torch.empty(B, Cin, Tz, dtype=z.dtype, device=z.device)
    blocks = _blocks(cfg, params, cond)
    _lib.call("gtk_decoder_reverse", ctypes.byref(d), blocks, _p(z), _p(mask), _p(x), _p(ws), ws_bytes, _stream())
    return x


@decoder_reverse.register_fake
def _(z, mask, params, cond, cfg, seed=None, energy=None, pitch=None):
    B, Cin, T = z.shape
    n = cfg[1]
    return z.new_empty((B, Cin, (T // n) * n))


@torch.library.custom_op(f"{NS}::decoder_backward", mutates_args=(), device_types="cuda")
def decoder_backward(dz: Tensor, dlogdet: Tensor, mask: Tensor, params: List[Tensor], cond: Optional[Tensor], ws: Tensor,
                     cfg: List[int], T: int, seed: Optional[Tensor] = None, energy: Optional[Tensor] = None,
                     pitch: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """Backward of decoder_forward(train=True) -> (dx [B,in,T], flat parameter grads (see grad_layout), dcond,
    d energy [B,T], d pitch [B,T]; empty tensors where the input was None)."""
    dz, dlogdet, mask = _chk(dz, "dz"), _chk(dlogdet, "dlogdet"), _chk(_fit_mask(mask, T), "mask")
    params = _chk_list(params, "params")
    cond = _opt(cond, "cond")
    B, Cin = dz.shape[0], dz.shape[1]
    energy, pitch = _track(energy, "energy", B, T), _track(pitch, "pitch", B, T)
    d_energy, d_pitch = _track_grads(dz, energy, pitch)
    d = _desc(cfg, B, T, seed, energy, pitch, d_energy if energy is not None else None, d_pitch if pitch is not None else None)
    dx = torch.empty(B, Cin, T, dtype=dz.dtype, device=dz.device)
    # one flat zero-initialised gradient buffer, viewed per parameter (one memset instead of ~430)
    offs, sizes, tot = grad_layout(params)
    flat = torch.zeros(tot, dtype=dz.dtype, device=dz.device)
    grads = [flat[o:o + s].view(p.shape) for o, s, p in zip(offs, sizes, params)]
    dcond = torch.zeros_like(cond) if cond is not None else dz.new_empty(0)
    blocks = _blocks(cfg, params, cond)
    gblocks = _blocks(cfg, grads, dcond if cond is not None else None)
    _lib.call("gtk_decoder_backward", ctypes.byref(d), blocks, gblocks, _p(mask), _p(dz), _p(dlogdet), _p(dx), _p(ws),
              ws.numel(), _stream())
    return dx, flat, dcond, d_energy, d_pitch


def _track_grads(like: Tensor, energy: Optional[Tensor], pitch: Optional[Tensor]) -> Tuple[Tensor, Tensor]:
    return (torch.empty_like(energy) if energy is not None else like.new_empty(0),
            torch.empty_like(pitch) if pitch is not None else like.new_empty(0))


def _track_fake(like, t, B, T):
    return like.new_empty((B, T)) if t is not None else like.new_empty(0)


@decoder_backward.register_fake
def _(dz, dlogdet, mask, params, cond, ws, cfg, T, seed=None, energy=None, pitch=None):
    B = dz.shape[0]
    return (dz.new_empty((B, dz.shape[1], T)), dz.new_empty((grad_layout(params)[2],)),
            torch.empty_like(cond) if cond is not None else dz.new_empty(0), _track_fake(dz, energy, B, T),
            _track_fake(dz, pitch, B, T))


_GRAD_SYNC = {"group": None, "segments": int(os.environ.get("GTK_GRAD_SYNC_SEGMENTS", "3"))}


def _world(group) -> int:
    import torch.distributed as dist
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1

# ---- flat-parameter variant: all decoder parameters live in ONE fp32 buffer (offsets in elements, 16-byte aligned),
# so autograd moves ONE gradient tensor instead of ~400 and the optimizer / NCCL see a single bucket.
def _flat_ptrs(flat: Tensor, offsets: Sequence[int]) -> List[int]:
    base = flat.data_ptr()
    return [base + 4 * int(o) for o in offsets]


@torch.library.custom_op(f"{NS}::decoder_forward_flat", mutates_args=(), device_types="cuda")
def decoder_forward_flat(x: Tensor, mask: Tensor, flat: Tensor, offsets: List[int], cond: Optional[Tensor], cfg: List[int],
                         train: bool, seed: Optional[Tensor] = None, energy: Optional[Tensor] = None,
                         pitch: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor]:
    """decoder_forward with the parameters given as one flat buffer + element offsets (same order as `params`)."""
    x = _chk(x, "x")
    B, Cin, T = x.shape
    mask = _chk(_fit_mask(mask, T), "mask")
    flat = _chk(flat, "flat")
    cond = _opt(cond, "cond")
    d = _desc(cfg, B, T, seed, _track(energy, "energy", B, T), _track(pitch, "pitch", B, T))
    lib = _lib.load()
    ws_bytes = lib.gtk_decoder_workspace_bytes(ctypes.byref(d), int(train))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    Tz = (T // d.n_sqz) * d.n_sqz
    z = torch.empty(B, Cin, Tz, dtype=x.dtype, device=x.device)
    logdet = torch.empty(B, dtype=x.dtype, device=x.device)
    blocks = _blocks(cfg, _flat_ptrs(flat, offsets), cond)
    _lib.call("gtk_decoder_forward", ctypes.byref(d), blocks, _p(x), _p(mask), _p(z), _p(logdet), _p(ws), ws_bytes,
              int(train), _stream())
    if not train:
        ws = ws.new_empty(0)
    return z, logdet, ws


@decoder_forward_flat.register_fake
def _(x, mask, flat, offsets, cond, cfg, train, seed=None, energy=None, pitch=None):
    B, Cin, T = x.shape
    n = cfg[1]
    return x.new_empty((B, Cin, (T // n) * n)), x.new_empty((B,)), x.new_empty((0,), dtype=torch.uint8)


@torch.library.custom_op(f"{NS}::decoder_backward_flat", mutates_args=(), device_types="cuda")
def decoder_backward_flat(dz: Tensor, dlogdet: Tensor, mask: Tensor, flat: Tensor, offsets: List[int], cond: Optional[Tensor],
                          ws: Tensor, cfg: List[int], T: int, seed: Optional[Tensor] = None, energy: Optional[Tensor] = None,
                          pitch: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """-> (dx, d flat (same layout as flat), dcond, d energy, d pitch)."""
    dz, dlogdet, mask = _chk(dz, "dz"), _chk(dlogdet, "dlogdet"), _chk(_fit_mask(mask, T), "mask")
    flat = _chk(flat, "flat")
    cond = _opt(cond, "cond")
    B, Cin = dz.shape[0], dz.shape[1]
    energy, pitch = _track(energy, "energy", B, T), _track(pitch, "pitch", B, T)
    d_energy, d_pitch = _track_grads(dz, energy, pitch)
    d = _desc(cfg, B, T, seed, energy, pitch, d_energy if energy is not None else None, d_pitch if pitch is not None else None)
    dx = torch.empty(B, Cin, T, dtype=dz.dtype, device=dz.device)
    dflat = torch.zeros_like(flat)
    dcond = torch.zeros_like(cond) if cond is not None else dz.new_empty(0)
    blocks = _blocks(cfg, _flat_ptrs(flat, offsets), cond)
    gblocks = _blocks(cfg, _flat_ptrs(dflat, offsets), dcond if cond is not None else None)
    _lib.call("gtk_decoder_backward", ctypes.byref(d), blocks, gblocks, _p(mask), _p(dz), _p(dlogdet), _p(dx), _p(ws),
              ws.numel(), _stream())
    return dx, dflat, dcond, d_energy, d_pitch


@decoder_backward_flat.register_fake
def _(dz, dlogdet, mask, flat, offsets, cond, ws, cfg, T, seed=None, energy=None, pitch=None):
    B = dz.shape[0]
    return (dz.new_empty((B, dz.shape[1], T)), torch.empty_like(flat),
            torch.empty_like(cond) if cond is not None else dz.new_empty(0), _track_fake(dz, energy, B, T),
            _track_fake(dz, pitch, B, T))


def range_bounds(nb: int, n_segments: int) -> List[int]:
    """Block index boundaries, ascending, of the backward ranges of a data-parallel step: range k covers blocks
    bounds[k] .. bounds[k + 1] - 1.  The ranges run from the last block down, so range 0 is the one whose collective nothing can
    hide: it gets a quarter of a share (12 blocks, 3 segments: blocks 11..6 | 5..1 | 0); GTK_GRAD_SYNC_LAST=k sets its size."""
    n_segments = max(1, min(int(n_segments), nb))
    if n_segments == 1:
        return [0, nb]
    last = max(1, nb // (4 * n_segments))
    if os.environ.get("GTK_GRAD_SYNC_LAST"):                  # blocks in the last (exposed) range, for sweeps
        last = max(1, min(nb - n_segments + 1, int(os.environ["GTK_GRAD_SYNC_LAST"])))
    return [0] + [last + (nb - last) * k // (n_segments - 1) for k in range(n_segments)]


_RANGE_STREAMS = {}


def _range_stream(device):
    """Helper stream per device from which the collective of a deferred backward range is issued."""
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    if key not in _RANGE_STREAMS:
        _RANGE_STREAMS[key] = torch.cuda.Stream(device=device)
    return _RANGE_STREAMS[key]


def decoder_backward_flat_overlapped(dz: Tensor, dlogdet: Tensor, mask: Tensor, flat: Tensor, offsets: List[int],
                                     cond: Optional[Tensor], ws: Tensor, cfg: List[int], T: int, seed: Optional[Tensor],
                                     group, n_segments: int = 3, energy: Optional[Tensor] = None,
                                     pitch: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """decoder_backward_flat for data-parallel training: the backward is issued as `n_segments` block ranges
    (gtk_decoder_backward_range) and the flat-gradient slice of each finished range is all-reduced (averaged, as DDP does:
    train_ms_emo_lang_pitch.py:166) asynchronously while the next range computes; only the last range's collective is
    exposed.  The slices are contiguous because the flat buffer is laid out block by block."""
    import torch.distributed as dist
    dz, dlogdet, mask = _chk(dz, "dz"), _chk(dlogdet, "dlogdet"), _chk(_fit_mask(mask, T), "mask")
    flat = _chk(flat, "flat")
    cond = _opt(cond, "cond")
    B, Cin = dz.shape[0], dz.shape[1]
    energy, pitch = _track(energy, "energy", B, T), _track(pitch, "pitch", B, T)
    d_energy, d_pitch = _track_grads(dz, energy, pitch)
    d = _desc(cfg, B, T, seed, energy, pitch, d_energy if energy is not None else None, d_pitch if pitch is not None else None)
    dx = torch.empty(B, Cin, T, dtype=dz.dtype, device=dz.device)
    dflat = torch.zeros_like(flat)
    dcond = torch.zeros_like(cond) if cond is not None else dz.new_empty(0)
    blocks = _blocks(cfg, _flat_ptrs(flat, offsets), cond)
    gblocks = _blocks(cfg, _flat_ptrs(dflat, offsets), dcond if cond is not None else None)
    nb, ppb = int(cfg[6]), params_per_block(int(cfg[5]), False, _prosody(cfg))
    sync = group is not None and _world(group) > 1         # group None: segmentation only (tests)
    world = dist.get_world_size(group) if sync else 1
    avg = sync and dist.get_backend(group) == "nccl"
    n_segments = max(1, min(int(n_segments), nb))
    bounds = range_bounds(nb, n_segments)
    works = []
    # Deferred join (GTK_RANGE_DEFER=0 disables): a range that is not the last one returns without joining the library's side
    # stream (its last weight-gradient launch + its reduce); the collective is issued from a helper stream that waits for the
    # caller's stream AND that side-stream work, so the next range's kernels never wait for a reduce.
    defer = os.environ.get("GTK_RANGE_DEFER", "1") != "0" and dz.is_cuda
    for k in range(n_segments - 1, -1, -1):                                   # ranges from the last block down
        lo, hi = bounds[k], bounds[k + 1] - 1
        deferred = defer and lo > 0
        _lib.call("gtk_decoder_backward_range_deferred" if deferred else "gtk_decoder_backward_range", ctypes.byref(d), blocks,
                  gblocks, _p(mask), _p(dz), _p(dlogdet), _p(dx), _p(ws), ws.numel(), hi, lo, _stream())
        if not sync:
            continue                      # segmentation only: the last range joins the side stream
        e0 = int(offsets[lo * ppb])
        e1 = int(offsets[(hi + 1) * ppb]) if hi + 1 < nb else flat.numel()
        if deferred:
            helper = _range_stream(dz.device)
            helper.wait_stream(torch.cuda.current_stream(dz.device))
            _lib.call("gtk_decoder_backward_join", ctypes.c_void_p(helper.cuda_stream))
            with torch.cuda.stream(helper):
                works.append(dist.all_reduce(dflat[e0:e1], op=dist.ReduceOp.AVG if avg else dist.ReduceOp.SUM, group=group,
                                             async_op=True))
            dflat.record_stream(helper)
            continue
        works.append(dist.all_reduce(dflat[e0:e1], op=dist.ReduceOp.AVG if avg else dist.ReduceOp.SUM, group=group,
                                     async_op=True))
    for w in works:
        w.wait()                  # stream-ordered for NCCL: the current stream waits, the host does not
    if sync and not avg:
        dflat.div_(world)
    return dx, dflat, dcond, d_energy, d_pitch


def _decf_setup(ctx, inputs, output):
    x, mask, flat, offsets, cond, cfg, train, seed, energy, pitch = inputs
    ctx.set_materialize_grads(False)       # the activation workspace output has no gradient: do not zero-fill GBs for it
    ctx.save_for_backward(mask, cond, output[2], flat, seed, energy, pitch)
    ctx.cfg, ctx.T, ctx.train, ctx.has_cond, ctx.offsets = list(cfg), x.shape[2], train, cond is not None, list(offsets)


def _decf_bwd(ctx, dz, dlogdet, dws):
    if not ctx.train:
        raise RuntimeError("glowtts_b200: decoder_forward was run with train=False; no activations were kept")
    mask, cond, ws, flat, seed, energy, pitch = ctx.saved_tensors
    B = mask.shape[0]
    if dz is None:
        Tz = (ctx.T // ctx.cfg[1]) * ctx.cfg[1]
        dz = torch.zeros(B, ctx.cfg[0], Tz, dtype=mask.dtype, device=mask.device)
    if dlogdet is None:
        dlogdet = torch.zeros(B, dtype=mask.dtype, device=mask.device)
    group = _GRAD_SYNC["group"]
    if group is not None and _world(group) > 1:
        # data-parallel training (train_ms_emo_lang_pitch.py:166 DDP): the flat gradient is all-reduced range by range,
        # overlapped with the rest of the backward
        dx, dflat, dcond, de, dp = decoder_backward_flat_overlapped(dz.contiguous(), dlogdet.contiguous(), mask, flat, ctx.offsets,
                                                                    cond, ws, ctx.cfg, ctx.T, seed, group,
                                                                    _GRAD_SYNC["segments"], energy, pitch)
    else:
        dx, dflat, dcond, de, dp = decoder_backward_flat(dz.contiguous(), dlogdet.contiguous(), mask, flat, ctx.offsets, cond, ws,
                                                         ctx.cfg, ctx.T, seed, energy, pitch)
    return (dx, None, dflat, None, (dcond if ctx.has_cond else None), None, None, None,
            _track_like(de, energy), _track_like(dp, pitch))




def _track_like(d: Tensor, t: Optional[Tensor]) -> Optional[Tensor]:
    """Gradient of a track in the shape autograd saw ([B,T'] or [B,1,T'], possibly longer than the T the decoder used)."""
    if t is None:
        return None
    out = d.new_zeros(t.shape)
    out.reshape(t.shape[0], t.shape[-1])[:, :d.shape[1]].copy_(d)
    return out


def set_grad_sync(group, segments: Optional[int] = None) -> None:
    """Enable (group = a torch.distributed process group, or dist.group.WORLD) or disable (None) the
    all-reduce of the decoder's flat gradient buffer inside the backward.  `segments` = number of block ranges the
    backward is split into so that each range's all-reduce overlaps the next range's kernels (default 3)."""
    _GRAD_SYNC["group"] = group
    if segments is not None:
        _GRAD_SYNC["segments"] = int(segments)


def _dec_setup(ctx, inputs, output):
    x, mask, params, cond, cfg, train, seed, energy, pitch = inputs
    ctx.set_materialize_grads(False)
    ctx.save_for_backward(mask, cond, output[2], seed, energy, pitch, *params)
    ctx.cfg, ctx.T, ctx.train, ctx.has_cond = list(cfg), x.shape[2], train, cond is not None


def _dec_bwd(ctx, dz, dlogdet, dws):
    if not ctx.train:
        raise RuntimeError("glowtts_b200: decoder_forward was run with train=False; no activations were kept")
    mask, cond, ws, seed, energy, pitch, *params = ctx.saved_tensors
    B = mask.shape[0]
    if dz is None:
        Tz = (ctx.T // ctx.cfg[1]) * ctx.cfg[1]
        dz = torch.zeros(B, ctx.cfg[0], Tz, dtype=mask.dtype, device=mask.device)
    if dlogdet is None:
        dlogdet = torch.zeros(B, dtype=mask.dtype, device=mask.device)
    dx, flat, dcond, de, dp = decoder_backward(dz.contiguous(), dlogdet.contiguous(), mask, list(params), cond, ws, ctx.cfg,
                                               ctx.T, seed, energy, pitch)
    if _GRAD_SYNC["group"] is not None:
        # data-parallel training (train_ms_emo_lang_pitch.py:166 DDP): ONE NCCL all-reduce over the flat
        # gradient buffer the backward kernels wrote, averaged over ranks as DDP does.
        from .distributed import sync_flat_grad
        sync_flat_grad(flat, _GRAD_SYNC["group"])
    offs, sizes, _ = grad_layout(params)
    grads = [flat[o:o + n].view(p.shape) for o, n, p in zip(offs, sizes, params)]
    return dx, None, grads, (dcond if ctx.has_cond else None), None, None, None, _track_like(de, energy), _track_like(dp, pitch)


decoder_forward.register_autograd(_dec_bwd, setup_context=_dec_setup)
decoder_forward_flat.register_autograd(_decf_bwd, setup_context=_decf_setup)


# =============================================================================== speaker conditioning
def _ptr_array(ts: Sequence[Optional[Tensor]]):
    arr = (ctypes.c_void_p * len(ts))()
    for i, t in enumerate(ts):
        arr[i] = None if t is None else t.data_ptr()
    return arr


@torch.library.custom_op(f"{NS}::cond_layers", mutates_args=(), device_types="cuda")
def cond_layers(g: Tensor, weights: List[Tensor], n_blocks: int, folded: bool) -> Tensor:
    """WN.cond_layer(g) (modules.py:148-149) of every coupling block in one launch.  g [B,G]; weights = per block
    (weight_g [N,1,1], weight_v [N,G,1], bias [N]) -- or (weight [N,G,1], bias [N]) when folded -- flattened.
    -> cond [n_blocks, B, N]."""
    g = _chk(g, "g")
    weights = _chk_list(weights, "cond weights")
    per = 2 if folded else 3
    B, G = g.shape
    wv = [weights[per * k + (0 if folded else 1)] for k in range(n_blocks)]
    wg = None if folded else [weights[per * k] for k in range(n_blocks)]
    bs = [weights[per * k + per - 1] for k in range(n_blocks)]
    N = wv[0].shape[0]
    cond = torch.empty(n_blocks, B, N, dtype=g.dtype, device=g.device)
    _lib.call("gtk_cond_layers", None if wg is None else _ptr_array(wg), _ptr_array(wv), _ptr_array(bs), n_blocks, _p(g), _p(cond),
              B, G, N, _stream())
    return cond


@cond_layers.register_fake
def _(g, weights, n_blocks, folded):
    return g.new_empty((n_blocks, g.shape[0], weights[1 if not folded else 0].shape[0]))


@torch.library.custom_op(f"{NS}::cond_layers_bwd", mutates_args=(), device_types="cuda")
def cond_layers_bwd(g: Tensor, weights: List[Tensor], n_blocks: int, folded: bool, dcond: Tensor) -> Tuple[Tensor, Tensor]:
    """-> (dg [B,G], flat gradient of `weights` in order, each padded to a multiple of 4 floats: ops.grad_layout)."""
    g, dcond = _chk(g, "g"), _chk(dcond, "dcond")
    weights = _chk_list(weights, "cond weights")
    per = 2 if folded else 3
    B, G = g.shape
    offs, sizes, tot = grad_layout(weights)
    flat = torch.empty(tot, dtype=g.dtype, device=g.device)
    views = [flat[o:o + n] for o, n in zip(offs, sizes)]
    wv = [weights[per * k + (0 if folded else 1)] for k in range(n_blocks)]
    wg = None if folded else [weights[per * k] for k in range(n_blocks)]
    dwv = [views[per * k + (0 if folded else 1)] for k in range(n_blocks)]
    dwg = None if folded else [views[per * k] for k in range(n_blocks)]
    dbs = [views[per * k + per - 1] for k in range(n_blocks)]
    N = wv[0].shape[0]
    dg = torch.zeros_like(g)
    _lib.call("gtk_cond_layers_bwd", None if wg is None else _ptr_array(wg), _ptr_array(wv), n_blocks, _p(g), _p(dcond),
              None if dwg is None else _ptr_array(dwg), _ptr_array(dwv), _ptr_array(dbs), _p(dg), B, G, N, _stream())
    return dg, flat


@cond_layers_bwd.register_fake
def _(g, weights, n_blocks, folded, dcond):
    return torch.empty_like(g), g.new_empty((grad_layout(weights)[2],))


def _cond_setup(ctx, inputs, output):
    g, weights, n_blocks, folded = inputs
    ctx.save_for_backward(g, *weights)
    ctx.n_blocks, ctx.folded = n_blocks, folded


def _cond_bwd(ctx, dcond):
    g, *weights = ctx.saved_tensors
    dg, flat = cond_layers_bwd(g, list(weights), ctx.n_blocks, ctx.folded, dcond.contiguous())
    group = _GRAD_SYNC["group"]
    if group is not None and _world(group) > 1:
        # data-parallel training: the conditioner's gradients are not part of the decoder's flat buffer, so they are
        # averaged here -- one collective for all blocks (DDP reduces every parameter: train_ms_emo_lang_pitch.py:166)
        from .distributed import sync_flat_grad
        sync_flat_grad(flat, group)
    offs, sizes, _ = grad_layout(weights)
    grads = [flat[o:o + n].view(w.shape) for o, n, w in zip(offs, sizes, weights)]
    return dg, grads, None, None


cond_layers.register_autograd(_cond_bwd, setup_context=_cond_setup)


# =============================================================================== coupling-block pieces
@torch.library.custom_op(f"{NS}::weightnorm_fwd", mutates_args=(), device_types="cuda")
def weightnorm_fwd(g: Tensor, v: Tensor) -> Tensor:
    """torch.nn.utils.weight_norm (dim=0): w[m] = g[m] v[m] / ||v[m]||  (modules.py:124,131,141)."""
    g, v = _chk(g, "weight_g"), _chk(v, "weight_v")
    M = v.shape[0]
    w = torch.empty_like(v)
    _lib.call("gtk_weightnorm_fwd", _p(g), _p(v), _p(w), None, M, v.numel() // M, _stream())
    return w


@weightnorm_fwd.register_fake
def _(g, v):
    return torch.empty_like(v)


@torch.library.custom_op(f"{NS}::weightnorm_bwd", mutates_args=(), device_types="cuda")
def weightnorm_bwd(g: Tensor, v: Tensor, dw: Tensor) -> Tuple[Tensor, Tensor]:
    g, v, dw = _chk(g, "weight_g"), _chk(v, "weight_v"), _chk(dw, "dw")
    M = v.shape[0]
    dg, dv = torch.empty_like(g), torch.empty_like(v)
    _lib.call("gtk_weightnorm_bwd", _p(g), _p(v), _p(dw), _p(dg), _p(dv), M, v.numel() // M, 0, _stream())
    return dg, dv


@weightnorm_bwd.register_fake
def _(g, v, dw):
    return torch.empty_like(g), torch.empty_like(v)


def _wn_setup(ctx, inputs, output):
    ctx.save_for_backward(*inputs)


def _wn_bwd(ctx, dw):
    g, v = ctx.saved_tensors
    return weightnorm_bwd(g, v, dw.contiguous())


weightnorm_fwd.register_autograd(_wn_bwd, setup_context=_wn_setup)


def _conv_raw(x: Tensor, w: Tensor, bias: Optional[Tensor], dil: int, transpose: bool) -> Tensor:
    B, Cin, T = x.shape
    M, Cw, K = w.shape
    if transpose:   # data gradient: contract over w's output channels
        assert Cin == M
        n_out, n_in = Cw, M
    else:
        assert Cin == Cw
        n_out, n_in = M, Cw
    wp = torch.empty(w.numel(), dtype=w.dtype, device=w.device)
    _lib.call("gtk_conv1d_pack_weight", _p(w), _p(wp), M, Cw, K, int(transpose), _stream())
    out = torch.empty(B, n_out, T, dtype=x.dtype, device=x.device)
    _lib.call("gtk_conv1d", _p(x), _p(wp), _p(bias), None, None, _p(out), B, n_in, n_out, T, K, dil, 0, _stream())
    return out


@torch.library.custom_op(f"{NS}::conv1d", mutates_args=(), device_types="cuda")
def conv1d(x: Tensor, w: Tensor, bias: Optional[Tensor], dilation: int) -> Tensor:
    """F.conv1d(x, w, bias, padding=(K-1)*dil/2, dilation=dil), K odd.  x [B,Cin,T], w [M,Cin,K]."""
    return _conv_raw(_chk(x, "x"), _chk(w, "weight"), _opt(bias, "bias"), dilation, False)


@conv1d.register_fake
def _(x, w, bias, dilation):
    return x.new_empty((x.shape[0], w.shape[0], x.shape[2]))


@torch.library.custom_op(f"{NS}::conv1d_bwd", mutates_args=(), device_types="cuda")
def conv1d_bwd(x: Tensor, w: Tensor, dout: Tensor, dilation: int) -> Tuple[Tensor, Tensor, Tensor]:
    """-> (dx, dw, dbias) of conv1d."""
    x, w, dout = _chk(x, "x"), _chk(w, "weight"), _chk(dout, "dout")
    B, Cin, T = x.shape
    M, _, K = w.shape
    dx = _conv_raw(dout, w, None, dilation, True)
    dw = torch.zeros_like(w)
    _lib.call("gtk_conv1d_wgrad", _p(dout), _p(x), None, _p(dw), B, Cin, M, T, K, dilation, 0, _stream())
    db = torch.zeros(M, dtype=x.dtype, device=x.device)
    _lib.call("gtk_channel_sum", _p(dout), _p(db), B, M, T, 0, _stream())
    return dx, dw, db


@conv1d_bwd.register_fake
def _(x, w, dout, dilation):
    return torch.empty_like(x), torch.empty_like(w), x.new_empty((w.shape[0],))


def _conv_setup(ctx, inputs, output):
    x, w, bias, dil = inputs
    ctx.save_for_backward(x, w)
    ctx.dil, ctx.has_bias = dil, bias is not None


def _conv_bwd(ctx, dout):
    x, w = ctx.saved_tensors
    dx, dw, db = conv1d_bwd(x, w, dout.contiguous(), ctx.dil)
    return dx, dw, (db if ctx.has_bias else None), None


conv1d.register_autograd(_conv_bwd, setup_context=_conv_setup)


@torch.library.custom_op(f"{NS}::coupling_affine", mutates_args=(), device_types="cuda")
def coupling_affine(x: Tensor, end_out: Tensor, mask: Optional[Tensor], reverse: bool, sigmoid_scale: bool) -> Tuple[Tensor, Tensor]:
    """attentions.py:172-183: end_out = [m | logs] -> (z = [x0 | affine(x1)], logdet [B] (zeros in reverse))."""
    x, end_out = _chk(x, "x"), _chk(end_out, "end_out")
    mask = _opt(mask, "mask")
    B, C, T = x.shape
    z = torch.empty_like(x)
    logdet = torch.zeros(B, dtype=x.dtype, device=x.device)
    _lib.call("gtk_coupling_affine", _p(x), _p(end_out), _p(mask), _p(z), _p(logdet), B, C, T, int(reverse),
              int(sigmoid_scale), _stream())
    return z, logdet


@coupling_affine.register_fake
def _(x, end_out, mask, reverse, sigmoid_scale):
    return torch.empty_like(x), x.new_empty((x.shape[0],))


@torch.library.custom_op(f"{NS}::coupling_affine_bwd", mutates_args=(), device_types="cuda")
def coupling_affine_bwd(x: Tensor, end_out: Tensor, mask: Optional[Tensor], dz: Tensor, dlogdet: Tensor,
                        sigmoid_scale: bool) -> Tuple[Tensor, Tensor]:
    """-> (d end_out, dx) of the forward affine (dx0 = dz0 passthrough, dx1 = dz1 e^logs mask)."""
    x, end_out, dz, dlogdet = _chk(x, "x"), _chk(end_out, "end_out"), _chk(dz, "dz"), _chk(dlogdet, "dlogdet")
    mask = _opt(mask, "mask")
    B, C, T = x.shape
    dend = torch.empty_like(end_out)
    dx = torch.empty_like(x)
    _lib.call("gtk_coupling_affine_bwd", _p(x), _p(end_out), _p(mask), _p(dz), _p(dlogdet), _p(dend), _p(dx), B, C, T,
              int(sigmoid_scale), _stream())
    dx[:, : C // 2].copy_(dz[:, : C // 2])
    return dend, dx


@coupling_affine_bwd.register_fake
def _(x, end_out, mask, dz, dlogdet, sigmoid_scale):
    return torch.empty_like(end_out), torch.empty_like(x)


def _aff_setup(ctx, inputs, output):
    x, end_out, mask, reverse, sigmoid_scale = inputs
    ctx.save_for_backward(x, end_out, mask)
    ctx.reverse, ctx.sigmoid_scale = reverse, sigmoid_scale


def _aff_bwd(ctx, dz, dlogdet):
    if ctx.reverse:
        raise RuntimeError("glowtts_b200: the reverse flow is inference-only (no backward)")
    x, end_out, mask = ctx.saved_tensors
    if dz is None:
        dz = torch.zeros_like(x)
    if dlogdet is None:
        dlogdet = torch.zeros(x.shape[0], dtype=x.dtype, device=x.device)
    dend, dx = coupling_affine_bwd(x, end_out, mask, dz.contiguous(), dlogdet.contiguous(), ctx.sigmoid_scale)
    return dx, dend, None, None, None


coupling_affine.register_autograd(_aff_bwd, setup_context=_aff_setup)


# =============================================================================== bf16 tensor-core conv (frames)
def _pad64(n: int) -> int:
    return (n + 63) // 64 * 64


def tc_frames(x: Tensor, c0: int = 0, nc: Optional[int] = None) -> Tensor:
    """x fp32 [B,C,T] -> bf16 frames [B,T,pad64(nc)] (gtk_tc_frames_from_tmajor)."""
    x = _chk(x, "x")
    B, C, T = x.shape
    nc = C - c0 if nc is None else nc
    ld = _pad64(nc)
    out = torch.empty(B, T, ld, dtype=torch.bfloat16, device=x.device)
    _lib.call("gtk_tc_frames_from_tmajor", _p(x), _p(out), B, C, T, c0, nc, ld, _stream())
    return out


def tc_pack_weight(w: Tensor, transpose: bool = False) -> Tensor:
    """w fp32 [M,Cin,K] -> bf16 [K, rows, pad64(cols)] operand of tc_conv1d (forward or data-gradient form)."""
    w = _chk(w, "weight")
    M, Cin, K = w.shape
    rows, cols = (Cin, _pad64(M)) if transpose else (M, _pad64(Cin))
    rows = (rows + 15) // 16 * 16
    out = torch.empty(K, rows, cols, dtype=torch.bfloat16, device=w.device)
    _lib.call("gtk_tc_pack_weight", _p(w), _p(out), M, Cin, K, int(transpose), rows, cols, _stream())
    return out


def tc_conv1d(frames: Tensor, wp: Tensor, bias: Optional[Tensor], dilation: int = 1, mask: Optional[Tensor] = None,
              want_f32: bool = True, want_bf16: bool = False):
    """Frame-major conv on the tensor cores: frames bf16 [B,T,ld], wp bf16 [K,N,ld] -> fp32 [B,T,N] and/or
    bf16 frames [B,T,pad64(N)]."""
    B, T, ld = frames.shape
    K, N, ldw = wp.shape
    assert ld == ldw and frames.dtype == torch.bfloat16 and wp.dtype == torch.bfloat16
    out32 = torch.empty(B, T, N, dtype=torch.float32, device=frames.device) if want_f32 else None
    ldb = _pad64(N)
    out16 = torch.zeros(B, T, ldb, dtype=torch.bfloat16, device=frames.device) if want_bf16 else None
    _lib.call("gtk_tc_conv1d", _p(frames), ld, _p(wp), _p(_opt(bias, "bias")), _p(_opt(mask, "mask")), B, T, K, dilation, N,
              _p(out32), N, _p(out16), ldb, 1 if mask is not None else 0, _stream())
    return out32, out16
