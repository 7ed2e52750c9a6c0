"""Secondary BASELINE.json configs on one B200 (bench.py keeps to configs[1] + the MAS legs):

  configs[2]  base_blank.json inference: reverse decoder flow z -> mel, batch sweep x T_mel sweep, no_grad, store_inverse()
  configs[3]  base_blank_ms.json multi-speaker train step (gin_channels=256, g = speaker embedding) on one GPU
  a-9 / f-3   the live fork's 3-WN coupling blocks (pitch + energy tracks) through the module-level path

    python profiles/config_sweeps.py > profiles/config_sweeps_r01.json

CUDA events, 3 warm-up calls, inputs resident in HBM (far larger than L2 from B*T >= 16k frames on).
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from glowtts_b200 import models  # noqa: E402
from glowtts_b200.hot_path import GraphedReverse, train_hot_path  # noqa: E402

DEV = torch.device("cuda", 0)


def decoder(gin=0, precision="bf16", p_dropout=0.05):
    torch.manual_seed(bench.SEED)
    c = bench.CFG
    dec = models.FlowSpecDecoder(c["in_channels"], c["hidden_channels"], c["kernel_size"], c["dilation_rate"], c["n_blocks"],
                                 c["n_layers"], p_dropout=p_dropout, n_split=c["n_split"], n_sqz=c["n_sqz"], gin_channels=gin,
                                 precision=precision)
    with torch.no_grad():
        for b in range(c["n_blocks"]):
            dec.flows[3 * b + 2].end.weight.normal_(0, 0.01)
    return dec.to(DEV)


def timed(fn, steps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def inference_sweep():
    dec = decoder(gin=0).eval()
    dec.store_inverse()
    rows = []
    gen = torch.Generator(device=DEV).manual_seed(1)
    with torch.no_grad():
        for T in (500, 1000, 2000):
            for B in (1, 4, 16, 64, 256):
                z = torch.randn(B, 80, T, device=DEV, generator=gen)
                mask = torch.ones(B, 1, T, device=DEV)
                ms = timed(lambda: dec(z, mask, reverse=True), 5 if B * T > 100000 else 10)
                flop = bench.FLOP_PER_SQ_FRAME_FWD * 12 * (B * T // 2)
                row = {"B": B, "T_mel": T, "ms": ms, "mel_frames_per_s": B * T / (ms * 1e-3),
                       "tflops": flop / (ms * 1e-3) / 1e12}
                if B * T <= 64000:                # launch-bound region: the same call replayed as one CUDA graph
                    rev = GraphedReverse(dec, z, mask)
                    row["ms_graphed"] = timed(lambda: rev(), 10)
                    row["launches"] = int(rev.launches)
                    del rev
                rows.append(row)
                del z, mask
    return rows


def multispeaker_step():
    dec = decoder(gin=256).train()
    opt = torch.optim.AdamW(dec.flat_parameters(), lr=1e-4, fused=True)
    gen = torch.Generator().manual_seed(bench.SEED)
    y, y_len, x_m, x_len = [t.to(DEV) for t in bench.synth_batch(bench.B_PER_GPU, gen, pin=False)]
    g = torch.randn(bench.B_PER_GPU, 256, 1, device=DEV)

    def step():
        opt.zero_grad(set_to_none=True)
        out = train_hot_path(dec, y, y_len, x_m, None, x_len, g=g)
        out.loss.backward()
        opt.step()

    ms = timed(step, 10)
    return {"workload": "configs[3] on one GPU: train step with gin_channels=256 speaker conditioning, B=32, eager launches",
            "ms_per_step": ms, "mel_frames_per_s": int(y_len.sum()) / (ms * 1e-3)}


def prosody_step():
    dec = decoder(gin=256, precision="fp32", p_dropout=0.0).train()
    gen = torch.Generator().manual_seed(bench.SEED)
    B = 8
    y, y_len, _, _ = [t.to(DEV) for t in bench.synth_batch(B, gen, pin=False)]
    mask = (torch.arange(bench.T_MEL, device=DEV)[None, :] < y_len[:, None]).float().unsqueeze(1)
    g = torch.randn(B, 256, 1, device=DEV)
    pitch = torch.randn(B, 1, bench.T_MEL, device=DEV)
    energy = torch.randn(B, 1, bench.T_MEL, device=DEV)

    def step():
        dec.zero_grad(set_to_none=True)
        z, ld = dec(y, mask, g=g, pitch=pitch, energy=energy)
        (z.pow(2).mean() - ld.mean()).backward()

    ms = timed(step, 3)
    return {"workload": f"live fork (base_blank_emo_lang_pitch.json decoder): 3 WaveNets per coupling block (wn, wn_energy, "
                        f"wn_pitch), fwd+bwd, B={B}, T_mel=860, module-level fp32 path (not the fused tcgen05 call)",
            "ms_per_step": ms, "mel_frames_per_s": int(y_len.sum()) / (ms * 1e-3)}


if __name__ == "__main__":
    out = {"inference_sweep": inference_sweep(), "multispeaker_train_step": multispeaker_step(), "prosody": prosody_step()}
    print(json.dumps(out, indent=1))
