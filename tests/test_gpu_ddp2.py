"""Two-GPU check of the data-parallel backward (train_ms_emo_lang_pitch.py:166 DDP): the flat gradient after the
range-by-range all-reduce -- with deferred range joins (gtk_decoder_backward_range_deferred + gtk_decoder_backward_join, the
default) and with joined ranges -- equals a single-call backward followed by ONE all-reduce.  Runs profiles/ddp_defer_check.py
under torchrun on two GPUs of the box; skipped on a single-GPU box (the world-size-2 host logic is covered on CPU with gloo
in test_distributed_gloo.py)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_deferred_range_all_reduce_matches_single_call_backward():
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533",
                          os.path.join(ROOT, "profiles", "ddp_defer_check.py")],
                         env=env, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "ddp_defer_check ok" in out.stdout, out.stdout[-2000:]
