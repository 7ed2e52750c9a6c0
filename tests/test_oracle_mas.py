"""Pin the MAS oracle (oracle/mas_oracle.c) against the reference: the committed golden
vectors (made by the reference's compiled core.pyx) and, when oracle/_ref is present,
the compiled reference itself on fresh random cases.  CPU only."""
import glob
import os

import numpy as np
import pytest

import helpers
from oracle import mas as omas


def _golden_files(golden_dir):
    return sorted(glob.glob(os.path.join(golden_dir, "mas_*.npz")))


def test_golden_present(golden_dir):
    assert len(_golden_files(golden_dir)) >= 6


@pytest.mark.parametrize("name", ["mas_small_ragged", "mas_ties", "mas_square", "mas_tx1", "mas_medium", "mas_wide"])
def test_c_port_matches_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    b, tx, ty, seed, ragged, quant = (int(v) for v in g["case"])
    value, mask, t_xs, t_ys = helpers.mas_case(b, tx, ty, seed, bool(ragged), bool(quant))
    assert (t_xs == g["t_xs"]).all() and (t_ys == g["t_ys"]).all()
    path = omas.maximum_path_np(value, mask, impl="c_port")
    assert (helpers.path_to_index(path) == g["index"]).all()
    assert (path.sum((1, 2)) == g["path_sum"]).all()


@pytest.mark.parametrize("seed", range(6))
def test_c_port_matches_compiled_reference(seed):
    if omas.ref_core() is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(100 + seed)
    b, tx, ty = int(rng.integers(1, 6)), int(rng.integers(1, 60)), int(rng.integers(60, 200))
    value, mask, _, _ = helpers.mas_case(b, tx, ty, 200 + seed, ragged=True, quantised=bool(seed % 2))
    a = omas.maximum_path_np(value, mask, impl="c_port")
    r = omas.maximum_path_np(value, mask, impl="ref")
    assert (a == r).all()
    assert (omas.maximum_path_np(value, mask, impl="c_port_omp") == r).all()


def test_path_properties():
    value, mask, t_xs, t_ys = helpers.mas_case(5, 40, 90, 7)
    path = omas.maximum_path_np(value, mask)
    for i in range(5):
        p = path[i]
        assert p[t_xs[i]:, :].sum() == 0 and p[:, t_ys[i]:].sum() == 0
        assert (p[:, : t_ys[i]].sum(0) == 1).all()            # one token per frame
        idx = p[:, : t_ys[i]].argmax(0)
        assert idx[0] == 0 and idx[-1] == t_xs[i] - 1          # starts/ends at the corners
        d = np.diff(idx)
        assert ((d == 0) | (d == 1)).all()                     # monotone, no skips
        assert (p[: t_xs[i]].sum(1) >= 1).all()                # surjective


def test_empty_utterance():
    value = np.ones((2, 3, 5), np.float32)
    mask = np.zeros((2, 3, 5), np.float32)
    mask[1, :2, :4] = 1
    path = omas.maximum_path_np(value, mask)
    assert path[0].sum() == 0 and path[1].sum() == 4
