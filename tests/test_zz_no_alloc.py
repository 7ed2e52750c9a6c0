"""gtk_mas_forward_static: the allocation-free MAS entry point (no counterpart in the reference).
Called through the C ABI and compared with the existing entry point gtk_mas_forward -- bit-exact,
the bar every MAS test here uses -- on two calls in a row that share the library's built-in scratch array.
Collected last on purpose (pytest -x must not lose other tests to a failure here)."""
import ctypes

import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


def _run(entry, value, t_xs, t_ys):
    from glowtts_b200 import _lib
    B, T_x, T_y = value.shape
    path = torch.full_like(value, 7.0)                  # outputs are caller-owned; poisoned to show they are fully written
    dur = torch.full((B, T_x), -5, dtype=torch.int32, device=value.device)
    idx = torch.full((B, T_y), -5, dtype=torch.int32, device=value.device)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    if entry == "static":
        _lib.call("gtk_mas_forward_static", P(value), P(t_xs), P(t_ys), P(path), P(dur), P(idx), B, T_x, T_y, st)
    else:
        lib = _lib.load()
        need = lib.gtk_mas_workspace_bytes(B, T_x, T_y)
        ws = torch.empty(max(need, 1), dtype=torch.uint8, device=value.device)
        _lib.call("gtk_mas_forward", P(value), P(t_xs), P(t_ys), P(path), P(dur), P(idx), B, T_x, T_y, P(ws), need, 0, st)
    torch.cuda.synchronize()
    return path.cpu().numpy(), dur.cpu().numpy(), idx.cpu().numpy()


def test_static_scratch_entry_matches_gtk_mas_forward_on_two_calls_in_a_row():
    from glowtts_b200 import _lib
    lib = _lib.load()
    assert lib.gtk_mas_static_scratch_bytes() == 8 << 20
    # two different shapes back to back: the second one is smaller, so it runs over the first call's stale scratch
    for (b, tx, ty, seed, quant) in ((32, 180, 860, 3, False), (5, 97, 333, 4, True)):
        value, mask, t_xs, t_ys = helpers.mas_case(b, tx, ty, seed, ragged=True, quantised=quant)
        v = torch.from_numpy(value).cuda()
        xs, ys = torch.from_numpy(t_xs).cuda(), torch.from_numpy(t_ys).cuda()
        free0 = torch.cuda.mem_get_info()[0]
        got = _run("static", v, xs, ys)
        want = _run("workspace", v, xs, ys)
        for g, w, name in zip(got, want, ("path", "durations", "index")):
            assert np.array_equal(g, w), name               # bit-exact, as every MAS test in this suite
        assert got[0].sum() == t_ys.sum()                   # one token per frame inside the lengths
        assert torch.cuda.mem_get_info()[0] <= free0 + (64 << 20)


def test_static_scratch_entry_rejects_shapes_beyond_its_limit():
    from glowtts_b200 import _lib
    dummy = torch.zeros(16, device="cuda")
    ti = torch.zeros(4, dtype=torch.int32, device="cuda")
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    # 512 x 2000 x ceil(400 / 32) x 4 B = 53 MB > 8 MiB: rejected before anything is launched or touched
    with pytest.raises(_lib.GtkError, match="exceeds the built-in scratch"):
        _lib.call("gtk_mas_forward_static", P(dummy), P(ti), P(ti), P(dummy), None, None, 512, 400, 2000,
                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
