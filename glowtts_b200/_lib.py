"""ctypes binding of libglowtts_b200.so, the C-ABI CUDA library (include/glowtts_b200.h).

The prototypes are read from the header itself, so the header stays the single source of
truth for the boundary: every function it declares must be exported by the library, and a
missing library or symbol is a hard error -- there is no CPU fallback on the product path.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import Dict, List, Tuple

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
HEADER = os.path.join(ROOT, "include", "glowtts_b200.h")
DEBUG_HEADER = os.path.join(ROOT, "include", "glowtts_b200_debug.h")   # instrumented builds only (-DGTK_INSTRUMENT)
LIB_PATH = os.environ.get("GTK_LIB") or os.path.join(HERE, "lib", "libglowtts_b200.so")   # GTK_LIB: diagnostic builds

GTK_MAX_LAYERS = 8
_f32p = ctypes.c_void_p


class DecoderDesc(ctypes.Structure):
    """gtk_decoder_desc (include/glowtts_b200.h)."""
    _fields_ = [(n, ctypes.c_int) for n in (
        "B", "in_channels", "T", "n_sqz", "hidden", "kernel_size", "dilation_rate", "n_layers",
        "n_blocks", "n_split", "sigmoid_scale", "precision")] + [("p_dropout", ctypes.c_float),
                                                                 ("seed", ctypes.c_ulonglong),
                                                                 ("seed_dev", ctypes.c_void_p),
                                                                 ("energy", ctypes.c_void_p), ("pitch", ctypes.c_void_p),
                                                                 ("d_energy", ctypes.c_void_p), ("d_pitch", ctypes.c_void_p)]


_BLOCK_FIELDS = ("an_logs", "an_bias", "ic_weight", "start_g", "start_v", "start_b",
                 "in_g", "in_v", "in_b", "rs_g", "rs_v", "rs_b", "end_w", "end_b", "cond")
_ARRAY_FIELDS = ("in_g", "in_v", "in_b", "rs_g", "rs_v", "rs_b")


class WnpParams(ctypes.Structure):
    """gtk_wnp_params / gtk_wnp_grads."""
    _fields_ = [(n, _f32p * GTK_MAX_LAYERS) for n in _ARRAY_FIELDS] + [(n, _f32p) for n in ("c1_g", "c1_v", "c1_b")]


class BlockParams(ctypes.Structure):
    """gtk_block_params; gtk_block_grads has the same member order and types."""
    _fields_ = [(n, (_f32p * GTK_MAX_LAYERS) if n in _ARRAY_FIELDS else _f32p) for n in _BLOCK_FIELDS] + [
        ("wn_energy", WnpParams), ("wn_pitch", WnpParams)]


BlockGrads = BlockParams

_CTYPES = {
    "int": ctypes.c_int, "size_t": ctypes.c_size_t, "float": ctypes.c_float,
    "void": None, "const char*": ctypes.c_char_p, "unsigned long long": ctypes.c_ulonglong,
    "long long": ctypes.c_longlong,
}

_PROTO = re.compile(r"^\s*(int|size_t|const char\*|unsigned long long)\s+(gtk_\w+)\s*\(([^;{]*)\)\s*;", re.M | re.S)


def parse_header(path: str = HEADER) -> Dict[str, Tuple[object, List[object]]]:
    """-> {name: (restype, [argtypes])} for every gtk_* prototype in the header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    protos = {}
    for ret, name, args in _PROTO.findall(text):
        argtypes = []
        args = " ".join(args.split())
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                else:
                    ty = a.rsplit(" ", 1)[0].replace("const ", "").strip()
                    argtypes.append(_CTYPES[ty])
        protos[name] = (_CTYPES[ret], argtypes)
    return protos


class GtkError(RuntimeError):
    pass


_lib = None
_protos = None


def load() -> ctypes.CDLL:
    """Load the CUDA library; raises if it has not been built (python -m glowtts_b200.build)."""
    global _lib, _protos
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GtkError(
            "glowtts_b200: %s is missing -- build it with `python -m glowtts_b200.build` "
            "(there is no CPU fallback)" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    protos = parse_header()
    for name, (restype, argtypes) in protos.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise GtkError("glowtts_b200: library does not export %s declared in %s" % (name, HEADER)) from e
        fn.restype = restype
        fn.argtypes = argtypes
    if hasattr(lib, "gtk_debug_wn_trace") and os.path.exists(DEBUG_HEADER):     # instrumented build: bind the probes too
        for name, (restype, argtypes) in parse_header(DEBUG_HEADER).items():
            fn = getattr(lib, name)
            fn.restype = restype
            fn.argtypes = argtypes
    _lib, _protos = lib, protos
    return lib


def declared_symbols() -> List[str]:
    return sorted(parse_header().keys())


def call(name: str, *args):
    """Call an int-returning entry point and raise GtkError with gtk_last_error() on failure."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.gtk_last_error()
        raise GtkError("%s failed (%d): %s" % (name, rc, msg.decode() if msg else ""))
    return rc


# ---- launch counter / per-kernel timing (used by bench.py) ----------------------------------------
def launch_count() -> int:
    return int(load().gtk_launch_count())


def prof_enable(on: bool) -> None:
    call("gtk_prof_reset")
    call("gtk_prof_enable", int(on))


def prof_read() -> Dict[str, Tuple[int, float]]:
    """-> {tag: (launches, total_ms)}; synchronises on the recorded events."""
    lib = load()
    out = {}
    for i in range(lib.gtk_prof_tag_count()):
        name = ctypes.create_string_buffer(64)
        n, ms = ctypes.c_int(0), ctypes.c_float(0.0)
        call("gtk_prof_read", i, name, 64, ctypes.byref(n), ctypes.byref(ms))
        if n.value:
            out[name.value.decode()] = (n.value, ms.value)
    return out
