#!/usr/bin/env python
"""bench.py -- the Glow-TTS train-step hot path on B200 (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one synthetic batch (configs/base.json sizes, B=32 per GPU):
decoder forward -> logp -> MAS -> prior gather -> mle_loss -> decoder backward (+ fused AdamW on the decoder
parameters; + NCCL all-reduce of the flat decoder gradient when N > 1).  One JSON line is printed by rank 0.

* value      : mel frames/s over all ranks with the batch already resident in HBM.
* e2e        : the same metric with the step's inputs copied from pinned host memory and the loss read back
               inside the timed region (through glowtts_b200.hot_path.train_hot_path, the user-facing call).
* roofline   : the dominant kernel family (the WaveNet convolutions), timed with CUDA events around every
               launch by the library's own profiler (gtk_prof_*) in a separate pass of the same steps.
* mas        : the MAS kernel alone on configs[0] (B=32, T_x=180, T_y=860) and its HBM roofline; mas_stress: the same
               on configs[4]'s per-GPU shape (B=512, T_x=400, T_y=2000); inference: the reverse decoder flow (configs[2]).
* cpu_baseline: the CPU restatement (oracle/) on a bounded sample, host cores of this box.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

# ---- workload: configs/base.json "model" section (reference configs/base.json:32-51) + BASELINE configs[1]
CFG = dict(in_channels=80, hidden_channels=192, kernel_size=5, dilation_rate=1, n_blocks=12, n_layers=4,
           n_split=4, n_sqz=2, gin_channels=0, p_dropout_dec=0.05)
B_PER_GPU, T_TEXT, T_MEL = 32, 180, 860
FLOP_PER_SQ_FRAME_FWD = 2 * (192 * 80 + 4 * (384 * 192 * 5) + 3 * (384 * 192) + 192 * 192 + 160 * 192)  # one block
SEED = 1234  # configs/base.json:5
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed `ncu --set full`
# capture (profiles/); bench.py never runs under a profiler, so the figure is carried here and updated per round.
NCU_TRAFFIC = {"mas configs[0]": 19.9e6,    # profiles/ncu_mas_cfg1_r01.csv: 19.89 MB read (value), 0 written (the 19.8 MB path stays in L2)
               "mas configs[4] on one GPU": 3.383e9,   # profiles/ncu_mas_cfg5_r01.csv: 1.728 GB read + 1.655 GB written (algorithmic 3.277 GB)
               "wn_layer_tc": 23.05e6,      # profiles/ncu_wn_layer_r01.txt: (16.89 + 0.82 MB, 27.50 + 0.89 MB) -> mean bytes / launch
               "wn_bwd_layer_tc": 24.9e6,   # profiles/ncu_wn_bwd_layer_r02.txt: 24.87 MB read + 0.01 MB written per launch (packed rows; r01: 33.0 MB)
               "wn_stack_tc": 54.1e6}       # profiles/ncu_wn_stack_r02.txt: 13.1 MB read + 41.0 MB written per launch (training saves)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=p["hbm_gbs"], tc_burst=p["bf16_tflops"], tc_sustained=p["bf16_tflops_sustained"], src="measured")
    except Exception:
        return dict(hbm=6650.0, tc_burst=1590.0, tc_sustained=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.thread = [], None, None
        self.index = index

    def start(self):
        # NVML from a polling thread (5 ms period: the timed region of a default run is ~0.1 s, too short for nvidia-smi's
        # loop mode to deliver more than a sample or two); nvidia-smi -lms as the fallback
        self.nvml_rows, self._stop = [], False
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid if not uuid.startswith("GPU-") else uuid).encode())
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            bits = {"hw_slowdown": pynvml.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": pynvml.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": pynvml.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": pynvml.nvmlClocksThrottleReasonSwPowerCap}

            def poll():
                while not self._stop:
                    try:
                        sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                        rs = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        self.nvml_rows.append((float(sm), float(mx), [n for n, b in bits.items() if rs & b]))
                    except Exception:
                        pass
                    time.sleep(0.005)
            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
            self.proc = "nvml"
            return
        except Exception:
            self.nvml_rows = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        def pump():
            for line in self.proc.stdout:
                self.rows.append(line.strip())
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        if self.proc == "nvml":
            self._stop = True
            self.thread.join(timeout=1)
            sm = sorted(r[0] for r in self.nvml_rows)
            reasons = sorted({n for r in self.nvml_rows for n in r[2]})
            med = sm[len(sm) * 3 // 4] if sm else None
            return {"sm_mhz": med, "sm_max_mhz": self.nvml_rows[0][1] if self.nvml_rows else None, "reasons": reasons,
                    "samples": len(sm), "source": "nvml, 5 ms polling during the timed region"}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        # median of the upper half = clocks under load (the sampler also sees the idle gaps)
        med = sm[len(sm) * 3 // 4] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def synth_batch(B: int, gen: torch.Generator, pin: bool = True, len_gen: torch.Generator = None):
    """LJSpeech-shaped synthetic batch (host, pinned).  `len_gen` draws the lengths (the same
    generator state on every rank, so all ranks do the same amount of work per step and weak-scaling efficiency by
    value equals efficiency by step time); `gen` draws the values (per-rank seed)."""
    len_gen = len_gen or gen
    y_len = (torch.randint(T_MEL // 2, T_MEL + 1, (B,), generator=len_gen) // 2) * 2
    y_len[0] = T_MEL
    x_len = torch.randint(T_TEXT // 2, T_TEXT + 1, (B,), generator=len_gen)
    x_len[0] = T_TEXT
    x_len = torch.minimum(x_len, y_len)
    y = (torch.randn(B, CFG["in_channels"], T_MEL, generator=gen) - 5.0)
    x_m = torch.randn(B, CFG["in_channels"], T_TEXT, generator=gen) * 0.5
    mask = (torch.arange(T_MEL)[None, :] < y_len[:, None]).float().unsqueeze(1)
    y = y * mask
    return [t.pin_memory() if pin else t for t in (y, y_len, x_m, x_len)]


def build_decoder(device, precision, gin=0):
    from glowtts_b200 import models
    torch.manual_seed(SEED)
    dec = models.FlowSpecDecoder(CFG["in_channels"], CFG["hidden_channels"], CFG["kernel_size"], CFG["dilation_rate"],
                                 CFG["n_blocks"], CFG["n_layers"], p_dropout=CFG["p_dropout_dec"], n_split=CFG["n_split"],
                                 n_sqz=CFG["n_sqz"], gin_channels=gin, precision=precision)
    with torch.no_grad():
        for b in range(CFG["n_blocks"]):          # non-trivial couplings (the reference starts them at zero)
            dec.flows[3 * b + 2].end.weight.normal_(0, 0.01)
    return dec.to(device)


class Runner:
    """Holds the resident model/optimizer and runs one step, device-resident or end to end."""

    def __init__(self, device, world, precision, graph=False, gin=0, prosody=False):
        from glowtts_b200 import ops
        self.device, self.world = device, world
        self.dec = build_decoder(device, precision, gin)
        # train_ms_emo_lang_pitch.py:126-130 AdamW; the decoder hands its weights over as one flat tensor
        self.opt = torch.optim.AdamW(self.dec.flat_parameters(prosody=prosody), lr=1e-4, fused=True, capturable=graph)
        self.graphed = None
        gen = torch.Generator().manual_seed(SEED + int(os.environ.get("RANK", "0")))
        self.host = synth_batch(B_PER_GPU, gen, len_gen=torch.Generator().manual_seed(SEED))
        if prosody:
            # the live fork's per-frame tracks (models.py:1057-1075): raw F0 in Hz with unvoiced frames at 0, raw energy
            pitch = torch.rand(B_PER_GPU, 1, T_MEL, generator=gen) * 300 + 60
            pitch = pitch * (torch.rand(B_PER_GPU, 1, T_MEL, generator=gen) > 0.3)
            energy = torch.rand(B_PER_GPU, 1, T_MEL, generator=gen) * 4 + 0.05
            self.host += [pitch.pin_memory(), energy.pin_memory()]
        self.dev = [t.to(device) for t in self.host]
        # configs[3] (base_blank_ms.json): speaker embedding g [B, gin, 1] (models.py:1040-1041), resident
        self.g = (torch.randn(B_PER_GPU, gin, 1, generator=gen).to(device) if gin else None)
        self.frames = int(self.host[1].sum())
        self.cells = int((self.host[1] * self.host[3]).sum())
        self.loss_host = torch.empty((), pin_memory=True)
        if world > 1:
            import torch.distributed as dist
            ops.set_grad_sync(dist.group.WORLD)
        if graph:
            from glowtts_b200.hot_path import GraphedTrainStep
            try:
                self.graphed = GraphedTrainStep(self.dec, self.opt, self.dev, g=self.g)
            except Exception as e:                          # e.g. a collective that cannot be captured: launch eagerly
                print(f"bench.py: CUDA-graph capture failed ({e!r}); falling back to eager launches", file=sys.stderr)
                self.graphed = None

    def step(self, batch):
        from glowtts_b200.hot_path import train_hot_path
        y, y_len, x_m, x_len, *tracks = batch
        pitch, energy = tracks if tracks else (None, None)
        self.opt.zero_grad(set_to_none=True)
        out = train_hot_path(self.dec, y, y_len, x_m, None, x_len, g=self.g, pitch=pitch, energy=energy)
        out.loss.backward()
        self.opt.step()
        return out.loss

    def step_resident(self):
        if self.graphed is not None:
            return self.graphed.step()                     # inputs already in the static device buffers
        return self.step(self.dev)

    def step_e2e(self):
        if self.graphed is not None:
            # input pipeline: this step consumes the batch whose H2D copy was started during the previous step and
            # starts the copy of the next one on the copy stream (one pinned-host -> device batch copy per step)
            if not self.graphed._staged:
                self.graphed.prefetch(self.host)
            loss = self.graphed.step()
            self.graphed.prefetch(self.host)
        else:
            batch = [t.to(self.device, non_blocking=True) for t in self.host]
            loss = self.step(batch)
        self.loss_host.copy_(loss.detach(), non_blocking=True)
        return loss

    @property
    def h2d_bytes(self):
        return sum(t.numel() * t.element_size() for t in self.host)


def barrier_sync(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, steps, world):
    """EXACTLY `steps` calls between barrier+synchronize pairs, CUDA events on the current stream,
    max over ranks -> ms total."""
    barrier_sync(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    barrier_sync(world)
    from glowtts_b200.distributed import max_over_ranks
    return max_over_ranks(e0.elapsed_time(e1), torch.device("cuda", torch.cuda.current_device()))


def kernel_spans():
    """Busy span (us) of the LAST launch of the whole-WaveNet forward kernel and of the fused layer backward inside the
    REPLAYED CUDA GRAPH of the step (first CTA past griddepcontrol.wait -> last CTA done), from the per-CTA globaltimer
    stamps of the -DGTK_INSTRUMENT library variant (profiles/_instr, built by __graft_entry__.build()).  The product library
    carries no stamps, so the measurement runs in a child process that loads the variant through GTK_LIB
    (profiles/ws_span_in_graph.py).  Unlike the event-bracketed eager pass this is the kernel as it runs in the timed
    region: PDL-overlapped prologue excluded, same clocks."""
    root = os.path.dirname(os.path.abspath(__file__))
    lib = os.path.join(root, "profiles", "_instr", "libglowtts_b200.so")
    if not os.path.exists(lib):
        return {}
    try:
        # GTK_WB_POST=0: the step's last fused-layer-backward launch is layer 0 of block 0, which in the timed step also
        # carries the d-`start` GEMM; the span is that of a plain layer launch, the one `flop_per_launch` describes
        env = dict(os.environ, GTK_LIB=lib, GTK_WB_POST="0")
        for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
            env.pop(k, None)
        out = subprocess.run([sys.executable, os.path.join(root, "profiles", "ws_span_in_graph.py"), "--json"], env=env,
                             capture_output=True, text=True, timeout=240, cwd=root)
        line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
        return json.loads(line)
    except Exception as e:                                  # diagnostic: never hide the main numbers, but say why
        print(f"bench.py: no in-graph kernel spans ({e!r}); is profiles/_instr older than the library? "
              "(__graft_entry__.build() rebuilds both)", file=sys.stderr)
        return {}


def mas_alone(device, steps=20, B=32, TX=T_TEXT, TY=T_MEL, n_buf=8, name="configs[0]", world=1, want_path=True):
    """configs[0]: maximum_path on synthetic fp32 logp, B=32, T_x=180, T_y=860, full lengths (rotating over enough
    distinct inputs to exceed the 126 MB L2 between repeats); configs[4] per GPU: B=512, T_x=400, T_y=2000 (one
    3.3 GB input: far larger than L2)."""
    from glowtts_b200 import _lib, ops
    gen = torch.Generator(device=device).manual_seed(SEED)
    vals = [torch.randn(B, TX, TY, device=device, generator=gen) for _ in range(n_buf)]
    t_xs = torch.full((B,), TX, dtype=torch.int32, device=device)
    t_ys = torch.full((B,), TY, dtype=torch.int32, device=device)
    call = (lambda v: ops.mas_path(v, t_xs, t_ys)) if want_path else (lambda v: ops.mas_index(v, t_xs, t_ys))
    for i in range(3):
        call(vals[i % n_buf])
    barrier_sync(world)
    _lib.prof_enable(True)
    for i in range(steps):
        call(vals[i % n_buf])
    torch.cuda.synchronize()
    prof = _lib.prof_read()
    _lib.prof_enable(False)
    n, ms = prof.get("mas", (0, 0.0))
    cells = B * TX * TY
    from glowtts_b200.distributed import max_over_ranks
    per = max_over_ranks(ms / max(n, 1) * 1e-3, device)          # the slowest rank sets the job's time
    pk = peaks()
    bpc = 8 if want_path else 4                                   # fp32 value read (+ fp32 path written: the drop-in)
    gbs = cells * bpc / per / 1e9 if per > 0 else 0.0
    del vals
    return {"workload": f"{name}: maximum_path B={B} per GPU, T_x={TX} T_y={TY} fp32, full lengths"
                        + ("" if want_path else "; index + durations only (no [B,T_x,T_y] path plane: SURVEY 8f-1 extension)"),
            "cells_per_s": world * cells / per if per else 0.0, "per_gpu_cells_per_s": cells / per if per else 0.0,
            "n_gpus": world, "sharding": "replicas only: utterances are independent, no collective",
            "us_per_call": per * 1e6, "launches_timed": n,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
                         "traffic": NCU_TRAFFIC.get("mas " + name) if (want_path and (B, TX, TY) in ((32, 180, 860), (512, 400, 2000))) else None,
                         "algorithmic_bytes_per_cell": bpc, "peak_source": pk["src"], "per": "GPU"},
            "latency_floor_note": f"T_y={TY} dependent DP columns per utterance; {B} CTAs on 148 SMs"}


def mas_fused_alone(device, steps=20, B=32, TX=T_TEXT, TY=T_MEL, n_buf=12):
    """north_star's fused variant at configs[0]'s shape: logp + MAS as ONE kernel (gtk_mas_fused_logp) against the
    two-kernel sequence gtk_logp -> gtk_mas_forward on the same z / x_m (rotating over n_buf input sets so that the
    operands and the 19.8 MB logp tensor do not sit in L2 between repeats).  Algorithmic HBM bytes per cell: fused =
    4 (path) + operands; unfused = 4 (logp write) + 4 (logp read) + 4 (path) + operands."""
    from glowtts_b200 import ops
    gen = torch.Generator(device=device).manual_seed(SEED)
    zs = [torch.randn(B, 80, TY, device=device, generator=gen) for _ in range(n_buf)]
    xs = [torch.randn(B, 80, TX, device=device, generator=gen) * 0.5 for _ in range(n_buf)]
    t_xs = torch.full((B,), TX, dtype=torch.int32, device=device)
    t_ys = torch.full((B,), TY, dtype=torch.int32, device=device)
    if not ops.mas_fused_supported(80, TX, TY, False):
        return {"error": "shape not covered by the fused kernel"}

    def unfused(i):
        return ops.mas_path(ops.logp(xs[i], None, zs[i]), t_xs, t_ys)

    def fused(i):
        return ops.mas_fused_logp(zs[i], xs[i], t_xs, t_ys)

    from glowtts_b200 import _lib
    out = {}
    for name, fn in (("logp_then_mas", unfused), ("fused", fused)):
        for i in range(3):
            fn(i % n_buf)
        torch.cuda.synchronize()
        _lib.prof_enable(True)                    # CUDA events around each kernel launch, on the launching stream
        for i in range(steps):
            fn(i % n_buf)
        torch.cuda.synchronize()
        prof = _lib.prof_read()
        _lib.prof_enable(False)
        per = {k: v[1] / max(v[0], 1) * 1e3 for k, v in prof.items()}
        out[name + "_us"] = sum(per.values())
        out[name + "_kernels_us"] = per
    ia = unfused(0)[2]
    ib = fused(0)[2]
    cells = B * TX * TY
    out.update({"workload": f"logp + maximum_path from z [B,80,{TY}] and x_m [B,80,{TX}], B={B}, mean_only, full lengths",
                "cells_per_s_fused": cells / (out["fused_us"] * 1e-6), "cells_per_s_logp_then_mas": cells / (out["logp_then_mas_us"] * 1e-6),
                "path_agreement": float((ia == ib).float().mean()),
                "timing": f"library profiler: CUDA events around each kernel launch, mean of {steps} launches"})
    return out


def inference_alone(run, steps=10, world=1):
    """configs[2] at the train-step batch: the reverse decoder flow z -> mel (models.py:1230) after store_inverse(),
    no_grad, B=32 x T_mel=860 resident in HBM; CUDA events over `steps` calls."""
    dec = run.dec
    y, y_len = run.dev[0], run.dev[1]
    mask = (torch.arange(T_MEL, device=y.device)[None, :] < y_len[:, None]).float().unsqueeze(1)
    was_training = dec.training
    dec.eval()
    with torch.no_grad():
        z = torch.randn_like(y) * mask
        for _ in range(3):
            dec(z, mask, reverse=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            dec(z, mask, reverse=True)
        e1.record()
        torch.cuda.synchronize()
    dec.train(was_training)
    from glowtts_b200.distributed import max_over_ranks
    ms = max_over_ranks(e0.elapsed_time(e1) / steps, y.device)
    frames = int(y_len.sum())
    flop = FLOP_PER_SQ_FRAME_FWD * CFG["n_blocks"] * (frames // CFG["n_sqz"])
    return {"workload": "reverse decoder flow z -> mel, B=32 per GPU, T_mel=860 (ragged), no_grad", "n_gpus": world,
            "mel_frames_per_s": world * frames / (ms * 1e-3), "ms_per_call": ms,
            "tflops_valid_frames_per_gpu": flop / (ms * 1e-3) / 1e12}


def multispeaker_step(device, world, precision, graph, steps, warmup):
    """BASELINE configs[3] (configs/base_blank_ms.json: gin_channels=256): the same train step with the speaker
    conditioner, batch sharded over the ranks, flat-gradient NCCL all-reduce; timed like the headline."""
    run = Runner(device, world, precision, graph=graph, gin=256)
    for _ in range(max(warmup, 3)):
        run.step_resident()
    ms = timed(run.step_resident, steps, world)
    from glowtts_b200.distributed import sum_over_ranks
    frames = sum_over_ranks(run.frames, device)
    out = {"workload": "configs[3]: base_blank_ms.json train step (gin_channels=256 speaker conditioning), B=32 per GPU",
           "n_gpus": world, "ms_per_step": ms / steps, "mel_frames_per_s": frames / (ms / steps * 1e-3),
           "launch": "cuda_graph" if run.graphed is not None else "eager"}
    del run
    torch.cuda.empty_cache()
    return out


def prosody_step(device, world, precision, graph, steps, warmup, ms_plain):
    """The live fork's train step: every coupling block chains wn -> wn_energy -> wn_pitch
    (attentions.py:155-160) inside the fused call, pitch / energy tracks copied with the batch; same B=32 per GPU.
    FLOPs: the three WaveNets of a block (3 x 4 layers) + start / end convs, forward, per squeezed frame."""
    run = Runner(device, world, precision, graph=graph, prosody=True)
    for _ in range(max(warmup, 3)):
        run.step_e2e()
    ms = timed(run.step_resident, steps, world)
    ms_e2e = timed(run.step_e2e, steps, world)
    from glowtts_b200 import _lib
    from glowtts_b200.distributed import sum_over_ranks
    frames = sum_over_ranks(run.frames, device)
    out = {"workload": "live-fork train step: 3 WaveNets per coupling block (wn, wn_energy, wn_pitch) with per-frame pitch / "
                       "energy tracks, B=32 per GPU, otherwise configs[1]",
           "n_gpus": world, "ms_per_step": ms / steps, "mel_frames_per_s": frames / (ms / steps * 1e-3),
           "e2e_ms_per_step": ms_e2e / steps, "h2d_bytes_per_step": run.h2d_bytes,
           "vs_one_wavenet_step": (ms / steps) / ms_plain if ms_plain else None,
           "launch": "cuda_graph" if run.graphed is not None else "eager",
           "gpu_launches_per_step": run.graphed.launches if run.graphed is not None else None}
    if world == 1:
        # roofline of the fused layer kernels in THIS step (event profiler, eager pass)
        _lib.prof_enable(True)
        for _ in range(2):
            run.step(run.dev)
        torch.cuda.synchronize()
        prof = _lib.prof_read()
        _lib.prof_enable(False)
        H, K, L = CFG["hidden_channels"], CFG["kernel_size"], CFG["n_layers"]
        flop_layer = (2.0 * (2 * H * H * K) + 2.0 * H * H * (2 * (L - 1) + 1) / L) * (run.frames // CFG["n_sqz"])
        pk = peaks()
        # the whole-WaveNet forward kernel (wn_stack.cu) runs all L layers of one WaveNet per launch
        for tag, fl in (("wn_stack_tc", flop_layer * L), ("wn_layer_tc", flop_layer), ("wn_bwd_layer_tc", flop_layer)):
            n_l, ms_l = prof.get(tag, (0, 0.0))
            if ms_l > 0:
                out[tag] = {"launches_per_step": n_l / 2, "ms_per_launch": ms_l / n_l,
                            "tflops": fl * n_l / (ms_l * 1e-3) / 1e12,
                            "frac_of_sustained_bf16_peak": fl * n_l / (ms_l * 1e-3) / 1e12 / pk["tc_sustained"]}
        fwd_flop_frame = CFG["n_blocks"] * (FLOP_PER_SQ_FRAME_FWD + 2 * 2 * (4 * (384 * 192 * 5) + 3 * (384 * 192) + 192 * 192))
        out["mflop_per_sq_frame_fwd"] = fwd_flop_frame / 1e6
        out["step_tflops"] = 3.0 * fwd_flop_frame * (run.frames // CFG["n_sqz"]) / (ms / steps * 1e-3) / 1e12
    del run
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------ CPU legs
class CpuArm:
    """The CPU side of the same hot path through the oracle (torch CPU fp32 restatement of the reference modules + the
    reference's compiled core.pyx for MAS, else the C port), with a resident checkpoint and an AdamW step, as the
    reference's train loop has (train_ms_emo_lang_pitch.py:126-130, 295-312)."""

    def __init__(self, threads: int, mas_impl: str):
        import helpers
        torch.set_num_threads(threads)
        self.threads, self.mas_impl = threads, mas_impl
        sd = helpers.make_state_dict(dict(CFG), SEED, end_std=0.01)
        self.sd = {k: v.requires_grad_(True) for k, v in sd.items() if ".wn_pitch." not in k and ".wn_energy." not in k}
        self.opt = torch.optim.AdamW(list(self.sd.values()), lr=1e-4)

    def step(self, b_sample: int):
        """One fwd + bwd + optimizer pass over `b_sample` utterances of the workload batch.
        -> (frames, cells, seconds, seconds spent in MAS)."""
        from oracle import flow as oflow
        from oracle import mas as omas
        gen = torch.Generator().manual_seed(SEED)
        y, y_len, x_m, x_len = synth_batch(B_PER_GPU, gen, pin=False, len_gen=torch.Generator().manual_seed(SEED))
        y, y_len, x_m, x_len = y[:b_sample], y_len[:b_sample], x_m[:b_sample], x_len[:b_sample]
        mask = (torch.arange(T_MEL)[None, :] < y_len[:, None]).float().unsqueeze(1)
        t0 = time.perf_counter()
        self.opt.zero_grad(set_to_none=True)
        z, logdet = oflow.decoder(self.sd, y, mask, hidden=CFG["hidden_channels"], dilation_rate=CFG["dilation_rate"],
                                  n_blocks=CFG["n_blocks"], n_layers=CFG["n_layers"], n_split=CFG["n_split"], n_sqz=CFG["n_sqz"],
                                  drop=lambda blk, i, v: torch.nn.functional.dropout(v, CFG["p_dropout_dec"]))
        with torch.no_grad():
            lp = oflow.logp(x_m, torch.zeros_like(x_m), z)
            x_mask = (torch.arange(T_TEXT)[None, :] < x_len[:, None]).float()
            amask = x_mask[:, :, None] * mask
            tm = time.perf_counter()
            attn = omas.maximum_path(lp, amask, impl=self.mas_impl)       # the reference wrapper's steps, CPU tensors
            t_mas = time.perf_counter() - tm
            z_m, z_logs = oflow.expand_prior(attn, x_m, torch.zeros_like(x_m))
        loss = oflow.mle_loss(z, z_m, z_logs, logdet, mask)
        loss.backward()
        self.opt.step()
        dt = time.perf_counter() - t0
        return int(y_len.sum()), int((y_len * x_len).sum()), dt, t_mas


def _cpu_arm():
    from oracle import mas as omas
    threads = os.cpu_count() or 1
    mas_impl = "ref" if omas.ref_core() is not None else "c_port"
    return CpuArm(threads, mas_impl), threads, mas_impl


def cpu_baseline(budget_s: float = 20.0):
    arm, threads, mas_impl = _cpu_arm()
    frames, cells, dt, t_mas = arm.step(1)                                    # warm-up + calibration
    b = max(1, min(B_PER_GPU, int(budget_s / max(dt, 1e-3))))
    frames, cells, dt, t_mas = arm.step(b)
    return {"value": frames / dt, "unit": "mel frames/s", "cores": threads,
            "kind": "port", "sample": f"{b} of {B_PER_GPU} utterances of the same synthetic batch, 1 fwd+bwd+AdamW pass "
                                      f"({dt:.1f} s); torch CPU fp32 restatement of the reference modules with "
                                      f"{threads} threads; MAS = {'the reference core.pyx compiled as shipped (serial)' if mas_impl == 'ref' else 'C port (serial)'}",
            "mas_cells_per_s": cells / t_mas if t_mas > 0 else None, "mas_cores": 1}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores.
    /root/reference cannot travel to the GPU box, so the decoder/logp leg is the oracle port of the reference's
    PyTorch modules (same torch CPU kernels) and MAS is the reference's own core.pyx compiled into oracle/_ref.
    The line carries ITS OWN config (fp32, eager CPU, the utterances it really ran per step)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm, threads, mas_impl = _cpu_arm()
    _, _, dt1, _ = arm.step(1)
    total = args.steps + args.warmup
    # the whole W + K run stays within ~4 minutes: the full 32-utterance batch when the box is fast enough for that
    b = max(1, min(B_PER_GPU, int(240.0 / max(total, 1) / max(dt1, 1e-3))))
    for _ in range(args.warmup):
        arm.step(b)
    t_sum, frames_sum = 0.0, 0
    for _ in range(args.steps):
        frames, cells, dt, t_mas = arm.step(b)
        t_sum += dt
        frames_sum += frames
    val = frames_sum / t_sum
    sample = (f"each step = {b} of {B_PER_GPU} utterances of the workload batch, fwd+bwd+AdamW, {threads} threads; "
              f"MAS = {'reference core.pyx as shipped (1 core)' if mas_impl == 'ref' else 'C port (1 core)'}")
    cfg = workload_config(1, "f32", "eager-cpu")
    cfg.update({"batch_per_step": b, "batch_per_gpu": None, "global_batch": b, "device": "host CPU",
                "parallelism": f"{threads} CPU threads (torch intra-op); no GPU",
                "l2": "n/a (CPU arm)", "layout": "padded [B, C, T] tensors, as the reference computes them",
                "kind": "oracle port of the reference modules + reference core.pyx"})
    print(json.dumps({
        "impl": "reference", "metric": "train_hot_path_mel_frames_per_s", "value": val, "unit": "mel frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_sum / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": val, "unit": "mel frames/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "mel frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(world, precision, launch="cuda_graph"):
    return {"workload": "Glow-TTS configs/base.json train-step hot path: FlowSpecDecoder(80,192,k5,12 blocks,4 layers,"
                        "n_sqz 2) fwd+bwd + logp + MAS + prior gather + mle_loss + AdamW; synthetic LJSpeech-shaped "
                        "mels, ragged lengths",
            "batch_per_gpu": B_PER_GPU, "global_batch": B_PER_GPU * world, "t_text": T_TEXT, "t_mel": T_MEL,
            "precision": precision, "p_dropout_dec": CFG["p_dropout_dec"], "launch": launch,
            "l2": "per-step working set (activation workspace > 1 GB) exceeds the 126 MB L2",
            "layout": "varlen-packed rows inside the decoder (no work on padding frames)",
            "parallelism": f"dp{world} (batch sharded, one flat-gradient NCCL all-reduce per step)" if world > 1 else "single GPU"}


# ------------------------------------------------------------------------------------------ main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("GTK_PRECISION", "bf16"), choices=["fp32", "bf16"],
                    help="bf16 = tcgen05 tensor-core WN path (north_star); fp32 = exact-parity CUDA-core path")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel from the host instead of replaying "
                    "the captured CUDA graph of the step (hot_path.GraphedTrainStep)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        run_reference(args)
        return

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE line, the JSON: anything a library writes to fd 1 (NCCL prints its version there)
    # goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)

    from glowtts_b200 import _lib
    _lib.load()
    run = Runner(device, world, args.precision, graph=not args.no_graph)

    for _ in range(args.warmup):
        run.step_e2e()
    barrier_sync(world)

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    n0 = _lib.launch_count()
    ms_res = timed(run.step_resident, args.steps, world)
    launches = _lib.launch_count() - n0
    if run.graphed is not None:                            # replays do not pass through the library's host counter
        launches = run.graphed.launches * args.steps
    spans = {}
    ms_e2e = timed(run.step_e2e, args.steps, world)
    clk = clocks.stop() if rank == 0 else None

    # roofline pass: same steps with the library's event profiler on
    # Each eager step is enqueued behind a 12 ms device-side spin, so the host is a whole step ahead of the GPU and the
    # kernels run back to back: an event pair then brackets the kernel's time on the device, not the host's launch latency
    # (with an empty stream the first event of a pair completes at once and the GPU idles until the launch call arrives).
    _lib.prof_enable(True)
    prof_steps = min(args.steps, 3)
    for _ in range(prof_steps):
        torch.cuda._sleep(int(12e-3 * 1.9e9))
        run.step(run.dev)                                  # eager launches: the event profiler brackets each one
    torch.cuda.synchronize()
    prof = _lib.prof_read()
    _lib.prof_enable(False)

    from glowtts_b200.distributed import sum_over_ranks
    frames_total, cells_total = sum_over_ranks(run.frames, device), sum_over_ranks(run.cells, device)
    run_frames, run_h2d, run_graphed = run.frames, run.h2d_bytes, run.graphed is not None
    run_rows = int((run.host[1] // CFG["n_sqz"] + (CFG["kernel_size"] // 2) * CFG["dilation_rate"] ** (CFG["n_layers"] - 1)).sum())

    # ---- secondary legs, at EVERY N (all ranks run them: the timings are reduced over the ranks).  MAS and the reverse
    # flow shard by utterance with no collective ("replicas only"), so these curves should be linear; configs[3] is the
    # multi-speaker train step with the gradient all-reduce.
    from glowtts_b200 import ops as _ops
    legs = {}

    def leg(name, fn):
        try:
            legs[name] = fn()
        except Exception as e:                                  # report, never hide the main numbers
            legs[name] = {"error": repr(e)}
        torch.cuda.empty_cache()

    leg("mas", lambda: mas_alone(device, world=world))
    if hasattr(_ops, "mas_index"):
        leg("mas_index_only", lambda: mas_alone(device, world=world, want_path=False))
    leg("inference", lambda: inference_alone(run, world=world))
    if world == 1:
        leg("mas_fused", lambda: mas_fused_alone(device))
    del run
    torch.cuda.empty_cache()
    b_stress = max(1, 512 // world)
    leg("mas_stress", lambda: mas_alone(device, steps=5, B=b_stress, TX=400, TY=2000, n_buf=1, world=world,
                                        name="configs[4] on one GPU" if world == 1 else f"configs[4]: B=512 sharded over {world} GPUs"))
    if hasattr(_ops, "mas_index"):
        leg("mas_stress_index_only", lambda: mas_alone(device, steps=5, B=b_stress, TX=400, TY=2000, n_buf=1, world=world,
                                                       want_path=False, name=f"configs[4]: B=512 over {world} GPU(s)"))
    leg("multispeaker", lambda: multispeaker_step(device, world, args.precision, not args.no_graph, args.steps, args.warmup))
    if args.precision == "bf16":
        leg("prosody", lambda: prosody_step(device, world, args.precision, not args.no_graph, args.steps, args.warmup,
                                            ms_res / args.steps))

    if rank == 0 and world == 1 and run_graphed and args.precision == "bf16":
        spans = kernel_spans()                             # child process with the instrumented library variant
    if rank == 0:
        pk = peaks()
        s_res, s_e2e = ms_res / args.steps * 1e-3, ms_e2e / args.steps * 1e-3
        # dominant kernel: the fused WaveNet layer (k=5 dilated conv GEMM + gate + res/skip 1x1 GEMM, wn_tc.cu);
        # fp32 path: the k=5 CUDA-core convolution.  Algorithmic flops = valid (unpadded) squeezed frames only.
        sq_frames_valid = run_frames // CFG["n_sqz"]
        # rows the tensor-core kernels really run: the decoder works on the varlen-packed batch (pack.cu): valid squeezed
        # frames + G = 2 gap rows per utterance, rounded up to whole 128-row tiles
        sq_frames_padded = (run_rows + 127) // 128 * 128
        H, K, L = CFG["hidden_channels"], CFG["kernel_size"], CFG["n_layers"]
        if "wn_layer_tc" in prof or "wn_stack_tc" in prof:
            # dominant = the fused WaveNet kernel that takes more of the step: forward (k=5 conv GEMM -> gate -> res/skip
            # GEMM; wn_stack.cu runs the start conv and all L layers of a WaveNet in one launch, wn_tc.cu one layer) or
            # backward (res/skip data gradient -> gate backward -> k=5 data gradient, one layer per launch): same GEMM shapes.
            tag = max(("wn_stack_tc", "wn_layer_tc", "wn_bwd_layer_tc"), key=lambda t: prof.get(t, (0, 0.0))[1])
            # per frame and layer: 2H x (H K) conv + RS x H res/skip, RS = 2H except the last layer of a block (H); mean over layers
            flop_per_frame = 2.0 * (2 * H * H * K) + 2.0 * H * H * (2 * (L - 1) + 1) / L
            if tag == "wn_stack_tc":                        # L layers + the start 1x1 conv (in_channels * n_sqz / 2 -> H)
                flop_per_frame = flop_per_frame * L + 2.0 * H * (CFG["in_channels"] * CFG["n_sqz"] // 2)
        else:
            tag = "conv_f32_kN"
            flop_per_frame = 2.0 * (2 * H * H * K)
        n_l, ms_l = prof.get(tag, (0, 0.0))
        flop_per_launch = flop_per_frame * sq_frames_valid
        ach = (flop_per_launch * n_l / (ms_l * 1e-3) / 1e12) if ms_l > 0 else 0.0
        roof = {"kernel": tag, "bound": "tensor", "achieved": ach, "peak": pk["tc_sustained"], "unit": "TFLOP/s",
                "frac": ach / pk["tc_sustained"], "traffic": NCU_TRAFFIC.get(tag), "peak_source": pk["src"] + " bf16 sustained",
                "launches_per_step": n_l / max(prof_steps, 1), "ms_per_launch": ms_l / max(n_l, 1),
                "share_of_step": (ms_l / max(prof_steps, 1)) / (s_res * 1e3),
                "flop_per_launch": flop_per_launch, "frames_valid": sq_frames_valid, "frames_computed": sq_frames_padded,
                "how": f"CUDA events around every launch (library profiler), separate eager pass of {prof_steps} steps, each enqueued "
                       "behind a 12 ms device-side spin so that the stream never runs dry (no host launch latency inside a pair)",
                "kernel_span_us_in_graph": spans.get(tag),
                "frac_by_span": (flop_per_launch / (spans[tag] * 1e-6) / 1e12 / pk["tc_sustained"]) if spans.get(tag) else None,
                "span_note": "busy span of the step's last launch of this kernel inside the replayed CUDA graph (first CTA past "
                             "griddepcontrol.wait -> last CTA done), per-CTA globaltimer stamps of the instrumented library "
                             "variant in a child process (with GTK_WB_POST=0: a plain layer launch, without the d-start GEMM that "
                             "layer 0 carries in the timed step); the event-bracketed figure above is the stand-alone launch, "
                             "prologue and launch latency included",
                "spans_us_in_graph": spans or None,
                "by_tag_ms_per_step": {k: v[1] / max(prof_steps, 1) for k, v in sorted(prof.items())}}
        step_flop = 3.0 * FLOP_PER_SQ_FRAME_FWD * CFG["n_blocks"] * sq_frames_padded
        out = {
            "metric": "train_hot_path_mel_frames_per_s", "value": frames_total / s_res, "unit": "mel frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": s_res * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if args.precision == "fp32" else "bf16 (fp32 accumulate)", "data": "synthetic",
            "config": workload_config(world, args.precision, "cuda_graph" if run_graphed else "eager"),
            "e2e": {"value": frames_total / s_e2e, "unit": "mel frames/s", "ms_per_step": s_e2e * 1e3,
                    "h2d_bytes_per_step": run_h2d, "d2h_bytes_per_step": 4,
                    "h2d": "pinned host -> device on a copy stream, double-buffered: the copy of step i+1's batch overlaps step i"
                           if run_graphed else "pinned host -> device on the compute stream before the step"},
            "gpu_launches": int(launches),
            "mas_cells_per_step": cells_total, "step_tflops": step_flop * world / s_res / 1e12,
            "roofline": roof, "clocks": clk,
        }
        out.update(legs)
        if world == 1 and not args.no_cpu_baseline:
            try:
                out["cpu_baseline"] = cpu_baseline()
            except Exception as e:  # the oracle is a checker; its absence must not hide the GPU numbers
                out["cpu_baseline"] = {"value": None, "error": repr(e)}
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    sys.stdout.flush()
    sys.stderr.flush()
    if world > 1:
        # Every collective of the run is done.  Tearing NCCL down while the captured step graph (which holds the
        # all-reduce) is alive blocked at N=2 on B200, so: rendezvous once through the c10d store (no NCCL), then leave
        # without destroy_process_group.
        torch.cuda.synchronize()
        try:
            import datetime
            import torch.distributed as dist
            store = dist.distributed_c10d._get_default_store()
            store.add("bench_done", 1)
            if rank == 0:
                deadline = time.time() + 60
                while store.add("bench_done", 0) < world and time.time() < deadline:
                    time.sleep(0.05)
            else:
                store.wait(["bench_rank0_out"], datetime.timedelta(seconds=60))
        except Exception:
            pass
        try:
            if rank == 0:
                store.set("bench_rank0_out", "1")
                time.sleep(0.2)
        except Exception:
            pass
        os._exit(0)


if __name__ == "__main__":
    main()
