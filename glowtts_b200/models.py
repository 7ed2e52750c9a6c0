"""Drop-in for models.FlowSpecDecoder (models.py:719-789) on the CUDA library.

The module tree (``flows.{3b}`` ActNorm, ``flows.{3b+1}`` InvConvNear, ``flows.{3b+2}``
CouplingBlock) and therefore the state_dict keys match the reference; ``forward`` hands the whole
stack to ONE library call (gtk_decoder_forward / _reverse; backward through gtk_decoder_backward).
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from . import attentions, modules, ops


class FlowSpecDecoder(nn.Module):
    def __init__(self, in_channels, hidden_channels, kernel_size, dilation_rate, n_blocks, n_layers, p_dropout=0.,
                 n_split=4, n_sqz=2, sigmoid_scale=False, gin_channels=0, emoin_channels=0, precision="fp32"):
        super().__init__()
        self.in_channels = in_channels
        self.hidden_channels = hidden_channels
        self.kernel_size = kernel_size
        self.dilation_rate = dilation_rate
        self.n_blocks = n_blocks
        self.n_layers = n_layers
        self.p_dropout = p_dropout
        self.n_split = n_split
        self.n_sqz = n_sqz
        self.sigmoid_scale = sigmoid_scale
        self.gin_channels = gin_channels
        self.emoin_channels = emoin_channels
        self.precision = precision      # "fp32" (exact-parity CUDA-core path) | "bf16" (tcgen05 path)
        self.flows = nn.ModuleList()
        for _ in range(n_blocks):
            self.flows.append(modules.ActNorm(channels=in_channels * n_sqz))
            self.flows.append(modules.InvConvNear(channels=in_channels * n_sqz, n_split=n_split))
            self.flows.append(attentions.CouplingBlock(
                in_channels * n_sqz, hidden_channels, kernel_size=kernel_size, dilation_rate=dilation_rate,
                n_layers=n_layers, gin_channels=gin_channels, emoin_channels=emoin_channels, p_dropout=p_dropout,
                sigmoid_scale=sigmoid_scale, n_sqz=n_sqz))

    # ---------------------------------------------------------------- fused call plumbing
    def _cfg(self, folded: bool, prosody: bool = False):
        # nn.Dropout(p_dropout) on the WN in_layer outputs (modules.py:120,153) is active in training mode only.
        drop = self.training and self.p_dropout > 0
        return [self.in_channels, self.n_sqz, self.hidden_channels, self.kernel_size, self.dilation_rate,
                self.n_layers, self.n_blocks, self.n_split, int(bool(self.sigmoid_scale)),
                {"fp32": 0, "bf16": 1}[self.precision], int(folded),
                int(round(self.p_dropout * 1e6)) if drop else 0, 0, int(prosody)]

    _SEED_STEP = 0x9E3779B97F4A7C15 - (1 << 64)      # odd 64-bit increment (as a signed int64)

    def _next_seed(self, device):
        """Per-call dropout seed as a one-element int64 DEVICE tensor.  The state starts from torch's CPU generator
        (torch.manual_seed reproducibility) and advances on the device, so no call synchronises the host and a
        captured CUDA graph of the step draws a fresh mask at every replay."""
        if not (self.training and self.p_dropout > 0):
            return None
        st = getattr(self, "_drop_state", None)
        if st is None or st.device != device:
            st = torch.tensor([ops.draw_seed()], dtype=torch.int64).to(device)
            self._drop_state = st
        seed = st.clone()
        st.add_(self._SEED_STEP)
        return seed

    def peek_dropout_seed(self, device=None) -> int:
        """The seed the next training forward will use (synchronises; for tests and debugging)."""
        st = getattr(self, "_drop_state", None)
        if st is None:
            dev = device or next(self.parameters()).device
            st = torch.tensor([ops.draw_seed()], dtype=torch.int64).to(dev)
            self._drop_state = st
        return int(st.item()) & 0xFFFFFFFFFFFFFFFF

    def _params(self, prosody: bool = False):
        out, folded = [], None
        for b in range(self.n_blocks):
            p, f = attentions.block_param_list(self.flows[3 * b], self.flows[3 * b + 1], self.flows[3 * b + 2], prosody)
            assert folded is None or folded == f
            folded = f
            out += p
        return out, bool(folded)

    def _cond(self, g: Optional[torch.Tensor]):
        if g is None or self.gin_channels == 0:
            return None
        # cond_layer(g) of all blocks (weight-norm fold, GEMM, bias) in ONE library launch, [NB, B, 2HL]
        layers = [self.flows[3 * b + 2].wn.cond_layer for b in range(self.n_blocks)]
        folded = layers[0].folded
        ws = []
        for l in layers:
            ws += [l.weight, l.bias] if folded else [l.weight_g, l.weight_v, l.bias]
        return ops.cond_layers(g.reshape(g.shape[0], -1).to(torch.float32), ws, self.n_blocks, folded)

    def forward(self, x, x_mask, g=None, emo=None, pitch=None, energy=None, reverse=False):
        prosody = pitch is not None or energy is not None
        if prosody and self.precision != "bf16":
            # live-fork configuration (three WaveNets per coupling block) on the exact-parity fp32 path: flow by flow
            # through the module-level ops; the bf16 path runs the chain inside the fused call (attentions.py:155-160)
            return self._forward_unfused(x, x_mask, g, reverse, pitch=pitch, energy=energy)
        for b in range(self.n_blocks):          # data-dependent init, if requested (init.py:17-23)
            if not self.flows[3 * b].initialized:
                return self._forward_unfused(x, x_mask, g, reverse, pitch=pitch, energy=energy)
        params, folded = self._params(prosody)
        mask2 = x_mask.reshape(x_mask.shape[0], x_mask.shape[-1]).to(torch.float32)
        cond = self._cond(g)
        cfg = self._cfg(folded, prosody)
        seed = self._next_seed(x.device)
        if reverse:
            with torch.no_grad():
                return ops.decoder_reverse(x, mask2, [p.detach() for p in params], cond, cfg, seed, energy, pitch), None
        train = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in params))
        if train and not folded and all(p.requires_grad for p in params):
            flat, offsets = self._ensure_flat(params, prosody)
            z, logdet, _ = ops.decoder_forward_flat(x, mask2, flat, offsets, cond, cfg, True, seed, energy, pitch)
        else:
            z, logdet, _ = ops.decoder_forward(x, mask2, params, cond, cfg, train, seed, energy, pitch)
        return z, logdet

    # ---------------------------------------------------------------- flat parameter storage (training)
    def _ensure_flat(self, params, prosody: bool = False):
        """Keep every decoder parameter as a view into ONE fp32 buffer (like DDP/FSDP flat buffers): the fused call
        takes the buffer, autograd returns ONE gradient tensor, and `p.grad` of the named parameters become views of
        it (no per-parameter copies).  Rebuilt lazily if the module was moved (.to()/.cuda()) or reloaded."""
        offs, sizes, total = ops.grad_layout(params)
        flat = getattr(self, "_flat", None)
        ok = flat is not None and flat.numel() == total and flat.device == params[0].device
        if ok:
            base = flat.data_ptr()
            ok = all(p.data_ptr() == base + 4 * o for p, o in zip(params, offs))
        if not ok:
            if flat is not None and getattr(self, "_keep_flat_grad", False):
                # an optimizer holds the old flat tensor: rebuilding it here would silently stop the decoder's training
                raise RuntimeError("glowtts_b200: the decoder was moved or its parameters were re-bound after "
                                   "flat_parameters() handed the flat buffer to an optimizer; build the optimizer again "
                                   "from flat_parameters() (or move the module before creating it)")
            with torch.no_grad():
                buf = torch.zeros(total, dtype=torch.float32, device=params[0].device)
                owners = self._param_owners(prosody)
                for p, o, n in zip(params, offs, sizes):
                    buf[o:o + n].copy_(p.detach().reshape(-1))
                    owners[p.data_ptr()].data = buf[o:o + n].view(owners[p.data_ptr()].shape)
            flat = buf.requires_grad_(True)
            flat.register_post_accumulate_grad_hook(self._scatter_flat_grad)
            self._flat, self._flat_offs, self._flat_sizes = flat, offs, sizes
            self._flat_params = [owners_p for owners_p in self._ordered_owners(prosody)]
            self._flat_prosody = prosody
        return self._flat, self._flat_offs

    def flat_parameters(self, prosody: bool = False):
        """Parameters for the optimizer with the fused decoder weights as ONE tensor: [flat buffer] + every parameter
        that is not part of it (e.g. the WN cond_layers; with prosody=False also the wn_energy / wn_pitch WaveNets, which
        then take no part in the fused call -- pass prosody=True when training with pitch / energy tracks on the bf16
        path).  The named parameters stay views of the buffer, so
        state_dict()/checkpoints are unchanged; an elementwise optimizer (Adam, AdamW, SGD) then runs one multi-tensor
        launch instead of ~40 and reads the gradient the backward kernels wrote in place (no scatter)."""
        params, folded = self._params(prosody)
        if folded:
            raise RuntimeError("glowtts_b200: weight norm was removed (store_inverse/remove_weight_norm); nothing to train")
        flat, _ = self._ensure_flat(params, prosody)
        self._keep_flat_grad = True
        inside = {id(p) for p in self._ordered_owners(prosody)}
        return [flat] + [p for p in self.parameters() if id(p) not in inside]

    def _ordered_owners(self, prosody: bool = False):
        """nn.Parameters in decoder_forward order (ActNorm logs/bias are [1,C,1] parameters passed as [C] views)."""
        out = []
        for b in range(self.n_blocks):
            an, ic, cb = self.flows[3 * b], self.flows[3 * b + 1], self.flows[3 * b + 2]
            out += [an.logs, an.bias, ic.weight, cb.start.weight_g, cb.start.weight_v, cb.start.bias]
            for grp in (cb.wn.in_layers, cb.wn.res_skip_layers):
                out += [l.weight_g for l in grp] + [l.weight_v for l in grp] + [l.bias for l in grp]
            out += [cb.end.weight, cb.end.bias]
            for wnp in ((cb.wn_energy, cb.wn_pitch) if prosody else ()):
                for grp in (wnp.in_layers, wnp.res_skip_layers):
                    out += [l.weight_g for l in grp] + [l.weight_v for l in grp] + [l.bias for l in grp]
                c1 = wnp.cond_layer1
                out += [c1.weight_g, c1.weight_v, c1.bias]
        return out

    def _param_owners(self, prosody: bool = False):
        return {p.data_ptr(): p for p in self._ordered_owners(prosody)}

    def _scatter_flat_grad(self, flat):
        if getattr(self, "_keep_flat_grad", False):
            # the optimizer owns the flat tensor (flat_parameters()); the named parameters still get `.grad` -- as views
            # of the flat gradient, no copies -- so that `model.parameters()`-based utilities of the reference's loop
            # (commons.clip_grad_value_, grad-norm logging, GradScaler inf checks) see all 12 flow blocks
            g = flat.grad
            for p, o, n in zip(self._flat_params, self._flat_offs, self._flat_sizes):
                p.grad = g[o:o + n].view(p.shape)
            return
        g = flat.grad
        flat.grad = None
        for p, o, n in zip(self._flat_params, self._flat_offs, self._flat_sizes):
            v = g[o:o + n].view(p.shape)
            if p.grad is None:
                p.grad = v
            else:
                p.grad.add_(v)

    def _forward_unfused(self, x, x_mask, g=None, reverse=False, pitch=None, energy=None):
        """Flow by flow through the module-level ops (used for the DDI pass, where ActNorm statistics must be
        taken between flows, and for the pitch / energy conditioned blocks of the live fork)."""
        from . import commons
        flows = list(self.flows) if not reverse else list(reversed(self.flows))
        logdet_tot = 0 if not reverse else None
        if self.n_sqz > 1:
            x, x_mask = commons.squeeze(x, x_mask, self.n_sqz)
        for f in flows:
            x, logdet = f(x, x_mask, g=g, pitch=pitch, energy=energy, reverse=reverse)
            if not reverse:
                logdet_tot = logdet_tot + logdet
        if self.n_sqz > 1:
            x, x_mask = commons.unsqueeze(x, x_mask, self.n_sqz)
        return x, logdet_tot

    def store_inverse(self):
        for f in self.flows:
            f.store_inverse()
