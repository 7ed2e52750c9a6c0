"""Drop-in for attentions.CouplingBlock (attentions.py:89-194) on the CUDA library."""
from __future__ import annotations

import torch
from torch import nn

from . import modules, ops


class CouplingBlock(nn.Module):
    """Affine coupling: start 1x1 -> WN -> end 1x1 -> z1 = (m + e^logs x1) mask, logdet = sum logs mask.

    The fork's extra pitch / energy WaveNets (``wn_energy``, ``wn_pitch``: attentions.py:114-117, applied in series
    after ``wn``, attentions.py:155-160) are modules.WNP: identity when their conditioning is None
    (modules.py:323-324), a WN with a time-varying conditioner otherwise."""

    def __init__(self, in_channels, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=0,
                 emoin_channels=0, p_dropout=0, sigmoid_scale=False, n_sqz=2):
        super().__init__()
        self.in_channels = in_channels
        self.hidden_channels = hidden_channels
        self.kernel_size = kernel_size
        self.dilation_rate = dilation_rate
        self.n_layers = n_layers
        self.gin_channels = gin_channels
        self.emoin_channels = emoin_channels
        self.p_dropout = p_dropout
        self.sigmoid_scale = sigmoid_scale
        self.start = modules.WeightNormConv1d(in_channels // 2, hidden_channels, 1)
        end = modules.WeightNormConv1d(hidden_channels, in_channels, 1, weight_norm=False)
        end.weight.data.zero_()     # attentions.py:107-110: the block starts as the identity
        end.bias.data.zero_()
        self.end = end
        # module order as the reference registers them (attentions.py:114-117), so state_dict() iterates alike
        self.wn_pitch = modules.WNP(hidden_channels, kernel_size, dilation_rate, n_layers, p_dropout, 1, n_sqz)
        self.wn_energy = modules.WNP(hidden_channels, kernel_size, dilation_rate, n_layers, p_dropout, 1, n_sqz)
        self.wn = modules.WN(in_channels, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels, p_dropout)

    def forward(self, x, x_mask=None, reverse=False, g=None, emo=None, pitch=None, energy=None, **kwargs):
        if pitch is not None and pitch.dim() == 2:
            pitch = pitch.unsqueeze(1)          # [B,T] -> [B,1,T] (attentions.py:137-141)
        if energy is not None and energy.dim() == 2:
            energy = energy.unsqueeze(1)
        m2 = modules._mask2d(x_mask)
        mask = 1 if x_mask is None else x_mask
        x_0 = x[:, : self.in_channels // 2].contiguous()
        h = self.start(x_0) * mask
        h = self.wn(h, x_mask, g)
        h = self.wn_energy(h, x_mask, energy)
        h = self.wn_pitch(h, x_mask, pitch)
        out = self.end(h)
        z, logdet = ops.coupling_affine(x, out, m2, reverse, bool(self.sigmoid_scale))
        return z, (None if reverse else logdet)

    def store_inverse(self):
        self.wn.remove_weight_norm()
        self.wn_energy.remove_weight_norm()
        self.wn_pitch.remove_weight_norm()


def _wnp_param_list(wnp, folded):
    """wn_energy / wn_pitch tensors in decoder_forward order (ops.py: after end_b when cfg[13] is set)."""
    out, c1 = [], wnp.cond_layer1
    if folded:
        for grp in (wnp.in_layers, wnp.res_skip_layers):
            out += [l.weight for l in grp]
            out += [l.bias for l in grp]
        out += [c1.weight, c1.bias]
    else:
        for grp in (wnp.in_layers, wnp.res_skip_layers):
            out += [l.weight_g for l in grp]
            out += [l.weight_v for l in grp]
            out += [l.bias for l in grp]
        out += [c1.weight_g, c1.weight_v, c1.bias]
    return out


def block_param_list(actnorm, invconv, coupling, prosody=False):
    """Tensors of one (ActNorm, InvConvNear, CouplingBlock) triple in decoder_forward order, and whether
    the WaveNet's weight norm has been folded (store_inverse).  prosody: the wn_energy / wn_pitch tensors follow."""
    wn = coupling.wn
    folded = wn.in_layers[0].folded
    out = [actnorm.logs.view(-1), actnorm.bias.view(-1), invconv.weight]
    st = coupling.start
    if folded:
        # start keeps its weight norm after store_inverse (attentions.py:188-194 folds only the WNs)
        out += [st.effective_weight().detach() if not st.folded else st.weight, st.bias]
        for grp in (wn.in_layers, wn.res_skip_layers):
            out += [l.weight for l in grp]
            out += [l.bias for l in grp]
    else:
        out += [st.weight_g, st.weight_v, st.bias]
        for grp in (wn.in_layers, wn.res_skip_layers):
            out += [l.weight_g for l in grp]
            out += [l.weight_v for l in grp]
            out += [l.bias for l in grp]
    out += [coupling.end.weight, coupling.end.bias]
    if prosody:
        out += _wnp_param_list(coupling.wn_energy, folded) + _wnp_param_list(coupling.wn_pitch, folded)
    return out, folded
