"""Drop-in for attentions.CouplingBlock (attentions.py:89-194) on the CUDA library."""
from __future__ import annotations

import torch
from torch import nn

from . import modules, ops


class CouplingBlock(nn.Module):
    """Affine coupling: start 1x1 -> WN -> end 1x1 -> z1 = (m + e^logs x1) mask, logdet = sum logs mask.

    The fork's extra pitch / energy WaveNets (attentions.py:114-117) are identity when their
    conditioning is None (modules.py:323-324), which is the path implemented here; passing
    ``pitch``/``energy`` raises (a "next" row)."""

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
        self.wn = modules.WN(in_channels, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels, p_dropout)

    def forward(self, x, x_mask=None, reverse=False, g=None, emo=None, pitch=None, energy=None, **kwargs):
        if pitch is not None or energy is not None:
            raise NotImplementedError("glowtts_b200.CouplingBlock: pitch/energy WaveNets are not built yet")
        m2 = modules._mask2d(x_mask)
        mask = 1 if x_mask is None else x_mask
        x_0 = x[:, : self.in_channels // 2].contiguous()
        h = self.start(x_0) * mask
        h = self.wn(h, x_mask, g)
        out = self.end(h)
        z, logdet = ops.coupling_affine(x, out, m2, reverse, bool(self.sigmoid_scale))
        return z, (None if reverse else logdet)

    def store_inverse(self):
        self.wn.remove_weight_norm()


def block_param_list(actnorm, invconv, coupling):
    """Tensors of one (ActNorm, InvConvNear, CouplingBlock) triple in decoder_forward order, and whether
    the WaveNet's weight norm has been folded (store_inverse)."""
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
    return out, folded
