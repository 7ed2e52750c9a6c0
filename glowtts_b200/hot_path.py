"""The training-step hot path of FlowGenerator.forward (models.py:1075-1119) + commons.mle_loss
(train_ms_emo_lang_pitch.py:295) as one host-side function over the library ops.

    z, logdet = decoder(y, z_mask, g)                      models.py:1075
    logp      = log N(z; x_m, exp(x_logs))                 models.py:1076-1082  (no_grad)
    attn      = maximum_path(logp, attn_mask)              models.py:1083       (no_grad)
    z_m,z_logs= attn^T x_m, attn^T x_logs                  models.py:1118-1119  (as a gather)
    loss      = mle_loss(z, z_m, z_logs, logdet, z_mask)   commons.py:28-33

The text encoder that produces x_m / x_logs is outside the path: they
arrive as tensors (requires_grad allowed; their gradients flow back through the gather).
"""
from __future__ import annotations

import copy
from typing import NamedTuple, Optional

import torch

from . import _lib, commons, ops


class HotPathOut(NamedTuple):
    loss: torch.Tensor
    z: torch.Tensor
    logdet: torch.Tensor
    attn: Optional[torch.Tensor]      # [B,T_x,T_y] 0/1 (None unless asked for)
    durations: torch.Tensor           # int32 [B,T_x]  (= attn.sum(2), models.py:1085)
    index: torch.Tensor               # int32 [B,T_y]  token per frame, -1 in the padding


def sequence_mask(length: torch.Tensor, max_length: int) -> torch.Tensor:
    """commons.py:82-86 -> bool [B, max_length]."""
    return torch.arange(max_length, device=length.device)[None, :] < length[:, None]


def normalize_track(track: torch.Tensor, T_y: int) -> torch.Tensor:
    """models.py:1057-1073: a raw pitch / energy track [B,1,T] (0 = unvoiced / silence) -> log(track), zeros kept zero,
    cut to T_y frames."""
    t = track.reshape(track.shape[0], track.shape[-1])[:, :T_y]
    zero = t == 0.0
    out = torch.log(torch.clamp(t, min=torch.finfo(t.dtype).tiny))
    return out.masked_fill(zero, 0.0).unsqueeze(1)


def train_hot_path(decoder, y, y_lengths, x_m, x_logs, x_lengths, g=None, return_attn=False, pitch=None,
                   energy=None) -> HotPathOut:
    """y [B,80,T_y] mel, y_lengths [B] (multiples of n_sqz, models.py:1052-1053), x_m / x_logs [B,80,T_x]
    (x_logs None = mean_only), x_lengths [B]; pitch / energy: raw per-frame tracks [B,1,T_y] of the live fork
    (models.py:1057-1075), None when the model is trained without them."""
    n = decoder.n_sqz
    T_y = (y.shape[2] // n) * n
    y = y[:, :, :T_y]
    y_lengths = torch.div(y_lengths, n, rounding_mode="floor") * n                      # models.py:1052 (preprocess), on device
    z_mask = sequence_mask(y_lengths, T_y).to(y.dtype).unsqueeze(1)                     # models.py:1054
    if pitch is not None or energy is not None:
        z, logdet = decoder(y, z_mask, g=g, pitch=None if pitch is None else normalize_track(pitch, T_y),
                            energy=None if energy is None else normalize_track(energy, T_y))
    else:
        z, logdet = decoder(y, z_mask, g=g)
    with torch.no_grad():
        lp = ops.logp(x_m.detach(), None if x_logs is None else x_logs.detach(), z.detach())
        t_xs = x_lengths.to(torch.int32)
        t_ys = y_lengths.to(torch.int32)
        # the kernel only reads cells inside [0,t_x) x [0,t_y), so value*mask (__init__.py:11) need not be materialised
        if return_attn:
            attn, dur, idx = ops.mas_path(lp, t_xs, t_ys)
        else:
            attn = None                                    # nobody reads the [B,T_x,T_y] float path: do not write it
            dur, idx = ops.mas_index(lp, t_xs, t_ys)
    z_m = ops.gather_tokens(x_m, idx, T_y)
    z_logs = None if x_logs is None else ops.gather_tokens(x_logs, idx, T_y)
    loss = commons.mle_loss(z, z_m, z_logs, logdet, z_mask)
    return HotPathOut(loss, z, logdet, attn if return_attn else None, dur, idx)


class GraphedTrainStep:
    """One training step of the hot path (forward, backward, optimizer) captured as a CUDA graph for fixed shapes.

    The step is ~330 kernel launches of 5-70 us each: replayed as one graph the host cost is a single launch, so
    the GPU never waits for the Python / driver launch path.  Everything the step needs is graph-safe by
    construction: no host synchronisation anywhere on the path, the dropout seed lives (and advances) on the
    device, the flat parameter / gradient buffers keep their addresses.  Inputs are copied into static device
    buffers before each replay (`step(batch)`), so any batch of the captured shape can be fed.

        stepper = GraphedTrainStep(decoder, optimizer, example_batch)   # optimizer: capturable (e.g. fused AdamW)
        loss = stepper.step((y, y_lengths, x_m, x_lengths))             # host (pinned) or device tensors

    Batches of another shape need another instance (bucket by padded length, as the reference's bucket sampler
    already does: data_utils.py DistributedBucketSampler).
    """

    def __init__(self, decoder, optimizer, example_batch, x_logs=None, g=None, warmup: int = 3):
        dev = next(decoder.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("glowtts_b200.GraphedTrainStep: the decoder must live on a CUDA device")
        self.decoder, self.optimizer = decoder, optimizer
        self.static = [t.to(dev, copy=True) for t in example_batch]
        self.x_logs = None if x_logs is None else x_logs.to(dev, copy=True)
        self.g = None if g is None else g.to(dev, copy=True)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        # Warm-up off the capture (allocations, flat buffers, seed state).  Constructing the stepper must not train the
        # model: parameters, optimizer state and the dropout seed state are snapshotted and restored around it, so step 1
        # of the stepper is step 1 of the run (exact resume, trajectories comparable with the reference's).
        with torch.cuda.stream(side):
            self._eager(dry=True)                          # materialises flat buffers / lazily created state
            snap_p = [p.detach().clone() for g in optimizer.param_groups for p in g["params"]]
            snap_o = copy.deepcopy(optimizer.state_dict())
            seed_state = getattr(decoder, "_drop_state", None)
            snap_s = None if seed_state is None else seed_state.clone()
            for _ in range(max(1, warmup)):
                self._eager()
            with torch.no_grad():
                for p, q in zip([p for g in optimizer.param_groups for p in g["params"]], snap_p):
                    p.copy_(q)
            cur = optimizer.state_dict()
            for k, st in snap_o["state"].items():          # restore IN PLACE: capturable optimizers keep device tensors
                for name, v in st.items():
                    tgt = cur["state"][k][name]
                    if torch.is_tensor(tgt):
                        tgt.copy_(v)
            if not snap_o["state"]:                        # the optimizer had no state yet: reset what warm-up created
                for st in optimizer.state.values():
                    for v in st.values():
                        if torch.is_tensor(v):
                            v.zero_()
            if snap_s is not None:                         # the dry pass drew one seed: rewind to the state before it
                decoder._drop_state.copy_(snap_s - decoder._SEED_STEP)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        self.optimizer.zero_grad(set_to_none=True)
        n0 = _lib.launch_count()
        with torch.cuda.graph(self.graph):
            self.out = self._eager()
        self.launches = _lib.launch_count() - n0           # library kernels per replay

    def _eager(self, dry: bool = False) -> HotPathOut:
        y, y_len, x_m, x_len, *tracks = self.static      # optionally followed by (pitch, energy) raw tracks [B,1,T_y]
        pitch, energy = tracks if tracks else (None, None)
        self.optimizer.zero_grad(set_to_none=True)
        out = train_hot_path(self.decoder, y, y_len, x_m, self.x_logs, x_len, self.g, pitch=pitch, energy=energy)
        out.loss.backward()
        if not dry:
            self.optimizer.step()
        return out

    def step(self, batch=None) -> torch.Tensor:
        """Copy `batch` (same shapes as the example) into the static buffers, replay, return the (static) loss tensor.
        With batch=None the step consumes the batch staged by prefetch(), if any, else re-runs the resident inputs."""
        if batch is not None:
            for dst, src in zip(self.static, batch):
                dst.copy_(src, non_blocking=True)
        elif self._staged:
            cur = torch.cuda.current_stream()
            cur.wait_event(self._staged_ready)             # the H2D copies of this batch (copy stream) have landed
            for dst, src in zip(self.static, self._stage):
                dst.copy_(src, non_blocking=True)          # device -> device, ~10 MB: microseconds
            self._stage_free.record(cur)
            self._staged = False
        self.graph.replay()
        return self.out.loss

    _staged = False

    def prefetch(self, batch) -> None:
        """Input pipeline: start the host -> device copy of the NEXT batch (pinned host tensors) on a copy stream into
        staging buffers, so that it overlaps the step that is running; the next step() consumes it.  This is the
        double buffering a DataLoader with pin_memory + non_blocking gives the reference (data_utils.py / train loop)."""
        if not hasattr(self, "_stage"):
            self._stage = [torch.empty_like(t) for t in self.static]
            self._copy_stream = torch.cuda.Stream(device=self.static[0].device)
            self._staged_ready = torch.cuda.Event()
            self._stage_free = torch.cuda.Event()
            self._stage_free.record(torch.cuda.current_stream())
        self._copy_stream.wait_event(self._stage_free)     # the previous staged batch has been consumed
        with torch.cuda.stream(self._copy_stream):
            for dst, src in zip(self._stage, batch):
                dst.copy_(src, non_blocking=True)
            self._staged_ready.record(self._copy_stream)
        self._staged = True


class GraphedReverse:
    """The inference hot path z -> mel (models.py:1230: decoder(z, mask, g, reverse=True)) captured as a CUDA graph for
    one (batch, T_mel) shape.  The reverse flow is ~100 short launches; below B*T_mel ~ 16k frames the host launch
    path, not the GPU, sets the latency (profiles/config_sweeps_r01.json), so replaying them as one graph is what makes
    small-batch synthesis fast.  Call store_inverse() on the decoder first (as the reference's inference does).

        rev = GraphedReverse(decoder, z_example, mask_example, g=None)
        mel = rev(z, mask)            # same shapes as the examples; returns the static output tensor
    """

    def __init__(self, decoder, z, mask, g=None, warmup: int = 2):
        if z.device.type != "cuda":
            raise RuntimeError("glowtts_b200.GraphedReverse: tensors must live on a CUDA device")
        self.decoder = decoder
        self.z, self.mask = z.clone(), mask.clone()
        self.g = None if g is None else g.clone()
        dev = z.device
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):
                decoder(self.z, self.mask, g=self.g, reverse=True)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.out, _ = decoder(self.z, self.mask, g=self.g, reverse=True)
        self.launches = _lib.launch_count() - n0

    def __call__(self, z=None, mask=None, g=None) -> torch.Tensor:
        if z is not None:
            self.z.copy_(z, non_blocking=True)
        if mask is not None:
            self.mask.copy_(mask, non_blocking=True)
        if g is not None and self.g is not None:
            self.g.copy_(g, non_blocking=True)
        self.graph.replay()
        return self.out
