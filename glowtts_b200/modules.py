"""Drop-ins for the flow building blocks of the reference's modules.py, on the CUDA library.

Same constructor arguments, forward protocol ``f(x, x_mask, reverse=False, **kw) -> (z, logdet)``
and state_dict key names as the reference, so reference checkpoints load.
"""
from __future__ import annotations

import math

import torch
from torch import nn

from . import ops


def _mask2d(x_mask):
    return None if x_mask is None else x_mask.detach().reshape(x_mask.shape[0], x_mask.shape[-1]).to(torch.float32)


def _x_len(x, m2):
    if m2 is None:
        return torch.full((x.shape[0],), float(x.shape[2]), dtype=torch.float32, device=x.device)
    return ops.mask_len(m2)


class WeightNormConv1d(nn.Module):
    """Parameter holder with the names torch.nn.utils.weight_norm(nn.Conv1d) gives
    (``weight_g``, ``weight_v``, ``bias``; plain ``weight`` after remove_weight_norm) and the same
    default initialisation; the convolution itself runs in the CUDA library."""

    def __init__(self, in_channels, out_channels, kernel_size=1, dilation=1, weight_norm=True):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.dilation = kernel_size, dilation
        w = torch.empty(out_channels, in_channels, kernel_size)
        nn.init.kaiming_uniform_(w, a=math.sqrt(5))
        bound = 1.0 / math.sqrt(in_channels * kernel_size)
        self.bias = nn.Parameter(torch.empty(out_channels).uniform_(-bound, bound))
        if weight_norm:
            self.weight_g = nn.Parameter(w.flatten(1).norm(dim=1).view(-1, 1, 1))
            self.weight_v = nn.Parameter(w)
        else:
            self.weight = nn.Parameter(w)

    @property
    def folded(self) -> bool:
        return "weight" in self._parameters

    def remove_weight_norm(self):
        if self.folded:
            return
        with torch.no_grad():
            v = self.weight_v
            w = self.weight_g * v / v.flatten(1).norm(dim=1).view(-1, 1, 1)
        del self.weight_g, self.weight_v
        self.weight = nn.Parameter(w)

    def effective_weight(self):
        return self.weight if self.folded else ops.weightnorm_fwd(self.weight_g, self.weight_v)

    def forward(self, x):
        return ops.conv1d(x, self.effective_weight(), self.bias, self.dilation)


class ActNorm(nn.Module):
    """modules.py:575-619."""

    def __init__(self, channels, ddi=False, **kwargs):
        super().__init__()
        self.channels = channels
        self.initialized = not ddi
        self.logs = nn.Parameter(torch.zeros(1, channels, 1))
        self.bias = nn.Parameter(torch.zeros(1, channels, 1))

    def forward(self, x, x_mask=None, reverse=False, **kwargs):
        m2 = _mask2d(x_mask)
        if not self.initialized:
            self.initialize(x, x_mask)
            self.initialized = True
        z = ops.flow_mix(x, m2, self.logs.view(-1), self.bias.view(-1), None, 1, reverse)
        if reverse:
            return z, None
        return z, _ActNormLogdet.apply(self.logs.view(-1), _x_len(x, m2))

    def store_inverse(self):
        pass

    def set_ddi(self, ddi):
        self.initialized = not ddi

    def initialize(self, x, x_mask):
        """Data-dependent init (modules.py:607-619); in a multi-GPU job all-reduce the stats first
        (see parallel.all_reduce_ddi_stats)."""
        with torch.no_grad():
            stats = ops.actnorm_ddi_stats(x, _mask2d(x_mask))
            stats = self._reduce_stats(stats)
            logs, bias = ops.actnorm_ddi_finalize(stats)
            self.logs.data.copy_(logs.view_as(self.logs))
            self.bias.data.copy_(bias.view_as(self.bias))

    @staticmethod
    def _reduce_stats(stats):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(stats)
        return stats


class _ActNormLogdet(torch.autograd.Function):
    """sum(logs) * x_len (modules.py:597); d/dlogs_c = sum_b dlogdet_b x_len_b."""

    @staticmethod
    def forward(ctx, logs, x_len):
        ctx.save_for_backward(x_len)
        ctx.C = logs.numel()
        return ops.flow_logdet_const(logs, None, 0.0, x_len)

    @staticmethod
    def backward(ctx, dlogdet):
        (x_len,) = ctx.saved_tensors
        dlogs, _ = ops.flow_logdet_const_bwd(dlogdet.contiguous(), x_len, ctx.C, None, 0.0)
        return dlogs, None


class InvConvNear(nn.Module):
    """modules.py:622-668: invertible n_split x n_split channel mix, no permute copies, W^-1 and
    logdet(W) computed on the device by the library."""

    def __init__(self, channels, n_split=4, no_jacobian=False, **kwargs):
        super().__init__()
        assert n_split % 2 == 0
        self.channels = channels
        self.n_split = n_split
        self.no_jacobian = no_jacobian
        w_init = torch.linalg.qr(torch.randn(n_split, n_split))[0]
        if torch.det(w_init) < 0:
            w_init[:, 0] = -w_init[:, 0]
        self.weight = nn.Parameter(w_init)

    def forward(self, x, x_mask=None, reverse=False, **kwargs):
        b, c, t = x.shape
        assert c % self.n_split == 0
        m2 = _mask2d(x_mask)
        if reverse:
            winv = self.weight_inv if hasattr(self, "weight_inv") else ops.invconv_prep(self.weight.detach())[0]
            return ops.flow_mix(x, m2, None, None, winv, self.n_split, False), None
        z = ops.flow_mix(x, m2, None, None, self.weight, self.n_split, False)
        if self.no_jacobian:
            return z, 0
        x_len = _x_len(x, m2)
        logdet = _InvConvLogdet.apply(self.weight, x_len, float(c // self.n_split))
        return z, logdet

    def store_inverse(self):
        w = self.weight.detach()
        # parameter preparation of a 4x4 (modules.py:667-668); on the device when the module lives there
        self.weight_inv = ops.invconv_prep(w)[0] if w.is_cuda else torch.inverse(w.float()).to(w.dtype)


class _InvConvLogdet(torch.autograd.Function):
    """logdet(W) * (C / n_split) * x_len (modules.py:658) and its gradient W^-T * (...)."""

    @staticmethod
    def forward(ctx, weight, x_len, group_count):
        winv, ld = ops.invconv_prep(weight)
        ctx.save_for_backward(winv, x_len)
        ctx.group_count = group_count
        return ops.flow_logdet_const(None, ld, group_count, x_len)

    @staticmethod
    def backward(ctx, dlogdet):
        winv, x_len = ctx.saved_tensors
        _, dW = ops.flow_logdet_const_bwd(dlogdet.contiguous(), x_len, 0, winv, ctx.group_count)
        return dW, None, None


class WN(nn.Module):
    """modules.py:105-179: non-causal WaveNet stack.  Standalone use composes the library's conv,
    gate and weight-norm ops; inside FlowSpecDecoder the whole stack runs in the fused decoder call."""

    def __init__(self, in_channels, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=0, p_dropout=0):
        super().__init__()
        assert kernel_size % 2 == 1
        assert hidden_channels % 2 == 0
        self.in_channels = in_channels
        self.hidden_channels = hidden_channels
        self.kernel_size = kernel_size
        self.dilation_rate = dilation_rate
        self.n_layers = n_layers
        self.gin_channels = gin_channels
        self.p_dropout = p_dropout
        self.in_layers = nn.ModuleList()
        self.res_skip_layers = nn.ModuleList()
        if gin_channels != 0:
            self.cond_layer = WeightNormConv1d(gin_channels, 2 * hidden_channels * n_layers, 1)
        for i in range(n_layers):
            self.in_layers.append(WeightNormConv1d(hidden_channels, 2 * hidden_channels, kernel_size,
                                                   dilation=dilation_rate ** i))
            rs = 2 * hidden_channels if i < n_layers - 1 else hidden_channels
            self.res_skip_layers.append(WeightNormConv1d(hidden_channels, rs, 1))

    def conditioning(self, g):
        """cond_layer(g) (modules.py:148-149): g [B,gin,1] -> [B, 2H*L] (a [2HL x gin] x [gin x B] GEMM)."""
        if g is None:
            return None
        l = self.cond_layer
        ws = [l.weight, l.bias] if l.folded else [l.weight_g, l.weight_v, l.bias]
        return ops.cond_layers(g.reshape(g.shape[0], -1).to(torch.float32), ws, 1, l.folded)[0]

    def forward(self, x, x_mask=None, g=None, **kwargs):
        H = self.hidden_channels
        drop = self.training and self.p_dropout > 0
        seed = ops.draw_seed() if drop else 0
        mask = 1 if x_mask is None else x_mask
        cond = self.conditioning(g)
        output = None
        for i in range(self.n_layers):
            x_in = self.in_layers[i](x)
            if drop:                          # modules.py:153, with the library's counter-based mask
                x_in = x_in * ops.dropout_scale(seed, 0, i, self.p_dropout, x_in.shape[0], x_in.shape[2], 2 * H, x_in.device)
            g_l = None if cond is None else cond[:, 2 * H * i: 2 * H * (i + 1)].contiguous()
            acts = ops.gate_fwd(x_in, g_l)
            rs = self.res_skip_layers[i](acts)
            if i < self.n_layers - 1:
                x = (x + rs[:, :H]) * mask
                skip = rs[:, H:]
            else:
                skip = rs
            output = skip if output is None else output + skip
        return output * mask

    def remove_weight_norm(self):
        if self.gin_channels != 0:
            self.cond_layer.remove_weight_norm()
        for l in self.in_layers:
            l.remove_weight_norm()
        for l in self.res_skip_layers:
            l.remove_weight_norm()


class WNP(nn.Module):
    """modules.py:272-362: the fork's pitch / energy WaveNet.  Same body as WN, but the conditioner is
    ``cond_layer1`` (1x1 conv, gin_channels1 -> 2*H*L/n_sqz) applied to the UN-squeezed [B,gin1,T_y] track and then
    squeezed to [B,2HL,T'] -- a time-varying conditioner; the module is the identity when ``g1`` is None
    (modules.py:323-324).  Composes the library's conv, squeeze, gate (cond_mode 2) and weight-norm ops."""

    def __init__(self, hidden_channels, kernel_size, dilation_rate, n_layers, p_dropout=0, gin_channels1=0, n_sqz=2):
        super().__init__()
        assert kernel_size % 2 == 1
        assert hidden_channels % 2 == 0
        self.hidden_channels = hidden_channels
        self.kernel_size = kernel_size
        self.dilation_rate = dilation_rate
        self.n_layers = n_layers
        self.p_dropout = p_dropout
        self.n_sqz = n_sqz
        self.gin_channels1 = gin_channels1
        self.in_layers = nn.ModuleList()
        self.res_skip_layers = nn.ModuleList()
        if gin_channels1 != 0:
            self.cond_layer1 = WeightNormConv1d(gin_channels1, 2 * hidden_channels * n_layers // n_sqz, 1)
        for i in range(n_layers):
            self.in_layers.append(WeightNormConv1d(hidden_channels, 2 * hidden_channels, kernel_size,
                                                   dilation=dilation_rate ** i))
            rs = 2 * hidden_channels if i < n_layers - 1 else hidden_channels
            self.res_skip_layers.append(WeightNormConv1d(hidden_channels, rs, 1))

    def conditioning(self, g1):
        """cond_layer1(g1) then WNP.squeeze (modules.py:320-322, 353-362): [B,gin1,T_y] -> [B,2HL,T_y/n_sqz]."""
        cond = self.cond_layer1(g1.to(torch.float32).contiguous())
        if self.n_sqz > 1:
            cond, _ = ops.squeeze_op(cond, None, self.n_sqz)
        return cond

    def forward(self, x, x_mask=None, g1=None, g2=None, **kwargs):
        if g1 is None:
            return x
        H = self.hidden_channels
        drop = self.training and self.p_dropout > 0
        seed = ops.draw_seed() if drop else 0
        mask = 1 if x_mask is None else x_mask
        cond = self.conditioning(g1)
        assert cond.shape[2] == x.shape[2], "pitch/energy track length must be n_sqz x the squeezed frame count"
        output = None
        for i in range(self.n_layers):
            x_in = self.in_layers[i](x)
            if drop:                          # modules.py:331, with the library's counter-based mask
                x_in = x_in * ops.dropout_scale(seed, 0, i, self.p_dropout, x_in.shape[0], x_in.shape[2], 2 * H, x_in.device)
            acts = ops.gate_fwd(x_in, cond[:, 2 * H * i: 2 * H * (i + 1)].contiguous())
            rs = self.res_skip_layers[i](acts)
            if i < self.n_layers - 1:
                x = (x + rs[:, :H]) * mask
                skip = rs[:, H:]
            else:
                skip = rs
            output = skip if output is None else output + skip
        return output * mask

    def remove_weight_norm(self):
        if self.gin_channels1 != 0:
            self.cond_layer1.remove_weight_norm()
        for l in self.in_layers:
            l.remove_weight_norm()
        for l in self.res_skip_layers:
            l.remove_weight_norm()
