"""Busy span of the whole-WaveNet kernel's LAST launch inside the replayed CUDA graph of the train step (per-CTA globaltimer
stamps of the -DGTK_INSTRUMENT build), next to the same launch run eagerly.
    GTK_LIB=profiles/_instr/libglowtts_b200.so python profiles/ws_span_in_graph.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from glowtts_b200 import _lib

JSON = "--json" in sys.argv
OUT = {}


def spans(tag):
    torch.cuda.synchronize()
    N, n = 24, 160
    buf = (ctypes.c_ulonglong * (n * N))()
    _lib.call("gtk_debug_ws_trace", buf, n)
    t = np.array(list(buf), dtype=np.int64).reshape(n, N)[:, :20]
    t = t[t[:, 0] > t[:, 0].max() - 400_000]
    t = t[(np.diff(t, axis=1) >= 0).all(axis=1)]
    per = (t[:, 19] - t[:, 1]) / 1e3
    OUT.setdefault(tag, {})["wn_stack_tc"] = float(t[:, 19].max() - t[:, 1].min()) / 1e3
    print(f"{tag}: {len(t)} CTAs; first post-wait stamp -> last end {(t[:, 19].max() - t[:, 1].min()) / 1e3:.1f} us; "
          f"per CTA (post-wait -> end) median {np.median(per):.1f} us; GEMM 1 of layer 1 median {np.median(t[:, 7] - t[:, 6]) / 1e3:.2f} us")
    nb = 8
    wb = (ctypes.c_ulonglong * (n * 7))()
    _lib.call("gtk_debug_wb_trace", wb, n)
    u = np.array(list(wb), dtype=np.int64).reshape(n, 7)
    u = u[u[:, 0] > u[:, 0].max() - 200_000]
    u = u[(np.diff(u, axis=1) >= 0).all(axis=1)]
    OUT[tag]["wn_bwd_layer_tc"] = float(u[:, 6].max() - u[:, 1].min()) / 1e3
    print(f"{tag}: fused layer backward, {len(u)} CTAs: post-wait -> last end {(u[:, 6].max() - u[:, 1].min()) / 1e3:.1f} us; GEMM b median {np.median(u[:, 4] - u[:, 3]) / 1e3:.2f} us")

if JSON:
    import contextlib, io, json
for graph in ((True,) if JSON else (False, True)):
    run = bench.Runner(torch.device("cuda", 0), 1, "bf16", graph=graph)
    for _ in range(6):
        run.step_resident()
    if JSON:
        with contextlib.redirect_stdout(io.StringIO()):
            spans("graph replay")
    else:
        spans("graph replay" if graph else "eager")
    del run
    torch.cuda.empty_cache()
if JSON:
    print(json.dumps(OUT["graph replay"]))
