"""Drop-ins for the hot-path helpers of the reference's commons.py, on the CUDA library."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops


def fused_add_tanh_sigmoid_multiply(input_a, input_b, n_channels=None):
    """commons.py:61-68: tanh((a+b)[:, :n]) * sigmoid((a+b)[:, n:]).  ``n_channels`` (the
    reference passes a 1-element IntTensor) must equal half of a's channels when given."""
    if n_channels is not None:
        n = int(n_channels[0]) if hasattr(n_channels, "__getitem__") else int(n_channels)
        assert 2 * n == input_a.shape[1], "n_channels must be half of input_a's channels"
    if input_b is not None and input_b.shape != input_a.shape and input_b.shape[-1] != 1:
        input_b = input_b.expand_as(input_a)
    return ops.gate_fwd(input_a, input_b)


def _mask2d(x_mask: Optional[torch.Tensor]):
    return None if x_mask is None else x_mask.reshape(x_mask.shape[0], x_mask.shape[-1])


def squeeze(x, x_mask=None, n_sqz=2):
    """commons.py:339-351 -> (x_sqz [b, c*n, t//n], mask [b, 1, t//n])."""
    out, m = ops.squeeze_op(x, _mask2d(x_mask), n_sqz)
    if x_mask is None:
        m = torch.ones_like(m)
    return out, m.detach().unsqueeze(1)      # masks carry no gradient


def unsqueeze(x, x_mask=None, n_sqz=2):
    """commons.py:354-364 -> (x_unsqz [b, c//n, t*n], mask [b, 1, t*n])."""
    out, m = ops.unsqueeze_op(x, _mask2d(x_mask), n_sqz)
    if x_mask is None:
        m = torch.ones_like(m)
    return out, m.detach().unsqueeze(1)      # masks carry no gradient


def mle_loss(z, m, logs, logdet, mask):
    """commons.py:28-33."""
    loss, _ = ops.mle_loss_fwd(z, m, logs, logdet, _mask2d(mask))
    return loss


def generate_path(duration, mask):
    """commons.py:127-143: duration [b, t_x], mask [b, t_x, t_y] -> path [b, t_x, t_y]."""
    return ops.generate_path(duration.to(torch.float32), mask.to(torch.float32)).to(mask.dtype)


def logp(x_m, x_logs, z):
    """The inline block models.py:1076-1082 as a function: -> [b, t_x, t_y]."""
    with torch.no_grad():
        return ops.logp(x_m, x_logs, z)


def expand_prior(x_m, x_logs, index):
    """models.py:1118-1119 (attn^T x_m, attn^T x_logs) as a gather by the per-frame token index."""
    t_y = index.shape[1]
    return ops.gather_tokens(x_m, index, t_y), ops.gather_tokens(x_logs, index, t_y)


def clip_grad_value_(parameters, clip_value, norm_type=2):
    """commons.py:320-336: total gradient norm (|g|^p summed over all parameters, then ^(1/p)) and, when
    ``clip_value`` is not None, an in-place clamp of every gradient to [-clip_value, clip_value].

    The reference syncs the host once per parameter (``.item()``, ~300 per step); here gradients that are
    neighbours in memory (the decoder's flat gradient buffer is ONE range) are merged, each range is one
    launch, and the norm stays on the device: the return value is a 0-d CUDA tensor (``float()`` it to get
    the reference's python float)."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    grads = [p.grad for p in parameters if p.grad is not None]
    norm_type = float(norm_type)
    if not grads:
        return torch.zeros((), device="cuda")
    for g in grads:
        if not (g.is_cuda and g.dtype == torch.float32 and g.is_contiguous()):
            raise RuntimeError("glowtts_b200: clip_grad_value_ needs contiguous CUDA fp32 gradients (there is no CPU path)")
    acc = torch.zeros(1, dtype=torch.float32, device=grads[0].device)
    spans = sorted(((g.data_ptr(), g.numel(), g) for g in grads if g.numel() > 0), key=lambda t: t[0])
    i = 0
    while i < len(spans):
        ptr, n, first = spans[i]
        j = i + 1
        while j < len(spans) and spans[j][0] == ptr + 4 * n and spans[j][2].untyped_storage().data_ptr() == first.untyped_storage().data_ptr():
            n += spans[j][1]
            j += 1
        if j == i + 1:
            rng = first.view(-1)
        else:                           # neighbours in one storage: a single flat view over all of them
            rng = first.new_empty(0).set_(first.untyped_storage(), (ptr - first.untyped_storage().data_ptr()) // 4, (n,), (1,))
        ops.clip_grad_value(rng, 0.0 if clip_value is None else float(clip_value), clip_value is not None, norm_type, acc)
        i = j
    return acc[0] ** (1.0 / norm_type)
