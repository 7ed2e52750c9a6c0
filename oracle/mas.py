"""oracle/mas.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Python access to the two MAS checkers:

* ``c_port``  : oracle/mas_oracle.c (our C restatement of core.pyx:9-45), via ctypes.
* ``ref_core``: the reference's own core.pyx compiled into oracle/_ref by
  oracle/build.py (``maximum_path_c``), when present.

``maximum_path(value, mask, impl=...)`` restates the reference wrapper
/root/reference/monotonic_align/__init__.py:6-21 around either checker (the
wrapper's Python source cannot travel to the GPU box; its steps are: value*mask,
float32 copy, zero int32 path, lengths from mask.sum(1)[:,0] / mask.sum(2)[:,0],
call the core, cast back to the input dtype).
"""
from __future__ import annotations

import ctypes
import importlib.util
import os

import numpy as np

from . import build as _build

_C = None
_REF = {}


def c_port():
    global _C
    if _C is None:
        path = _build.oracle_lib_path()
        if not os.path.exists(path):
            _build.build_oracle()
        lib = ctypes.CDLL(path)
        i32p = ctypes.POINTER(ctypes.c_int32)
        f32p = ctypes.POINTER(ctypes.c_float)
        for name in ("mas_oracle_batch", "mas_oracle_batch_omp"):
            fn = getattr(lib, name)
            fn.argtypes = [i32p, f32p, i32p, i32p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float]
            fn.restype = None
        lib.mas_oracle_wrapper_prep.argtypes = [f32p, f32p, i32p, i32p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        lib.mas_oracle_wrapper_prep.restype = None
        _C = lib
    return _C


def ref_core(omp: bool = False):
    """The reference's compiled Cython module, or None if oracle/_ref was not built."""
    key = bool(omp)
    if key not in _REF:
        path = _build.ref_core_path()
        if omp:
            path = os.path.join(os.path.dirname(path), "omp", os.path.basename(path))
        if not os.path.exists(path) and os.path.exists(_build.REF_PYX):
            _build.build_ref()
        if not os.path.exists(path):
            _REF[key] = None
        else:
            spec = importlib.util.spec_from_file_location("monotonic_align.core", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            _REF[key] = mod
    return _REF[key]


def _ptr(a, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct))


def maximum_path_c_port(paths, values, t_xs, t_ys, max_neg_val=-1e9, omp=False):
    """core.pyx:40-45 signature, executed by the C port (mutates values, fills paths)."""
    assert paths.dtype == np.int32 and values.dtype == np.float32
    assert paths.flags.c_contiguous and values.flags.c_contiguous
    b, tx, ty = values.shape
    t_xs = np.ascontiguousarray(t_xs, dtype=np.int32)
    t_ys = np.ascontiguousarray(t_ys, dtype=np.int32)
    fn = c_port().mas_oracle_batch_omp if omp else c_port().mas_oracle_batch
    fn(_ptr(paths, ctypes.c_int32), _ptr(values, ctypes.c_float), _ptr(t_xs, ctypes.c_int32),
       _ptr(t_ys, ctypes.c_int32), b, tx, ty, ctypes.c_float(max_neg_val))


def maximum_path_np(value, mask, impl="c_port"):
    """numpy-in / numpy-out restatement of monotonic_align/__init__.py:11-20."""
    value = np.asarray(value) * np.asarray(mask)
    value = np.ascontiguousarray(value.astype(np.float32))
    path = np.zeros_like(value).astype(np.int32)
    mask = np.asarray(mask)
    t_x_max = mask.sum(1)[:, 0].astype(np.int32)
    t_y_max = mask.sum(2)[:, 0].astype(np.int32)
    if impl == "c_port":
        maximum_path_c_port(path, value, t_x_max, t_y_max)
    elif impl == "c_port_omp":
        maximum_path_c_port(path, value, t_x_max, t_y_max, omp=True)
    elif impl in ("ref", "ref_omp"):
        mod = ref_core(omp=(impl == "ref_omp"))
        if mod is None:
            raise RuntimeError("oracle/_ref not built (needs /root/reference at build time)")
        mod.maximum_path_c(path, value, np.ascontiguousarray(t_x_max), np.ascontiguousarray(t_y_max))
    else:
        raise ValueError(impl)
    return path


def maximum_path(value, mask, impl="c_port"):
    """torch-in / torch-out restatement of monotonic_align/__init__.py:6-21."""
    import torch

    device, dtype = value.device, value.dtype
    path = maximum_path_np((value * mask).data.cpu().numpy().astype(np.float32),
                           mask.data.cpu().numpy(), impl=impl)
    return torch.from_numpy(path).to(device=device, dtype=dtype)
