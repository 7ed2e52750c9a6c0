"""Import the reference's Python modules from /root/reference (this container only).

librosa is absent here and is imported at module top by commons.py:7 ->
audio_processing.py / stft.py, none of which is on the hot path, so stub modules are
installed first.  The reference's compiled Cython core is made importable as
``monotonic_align.monotonic_align.core`` is NOT needed: tests call oracle.mas.ref_core().
Returns None when /root/reference is absent (GPU box) -- callers skip.
"""
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"
_cached = None


def available() -> bool:
    return os.path.isdir(REFERENCE_ROOT)


def load():
    """-> namespace with .commons .modules .attentions .models, or None."""
    global _cached
    if _cached is not None or not available():
        return _cached
    if "librosa" not in sys.modules:
        lib = types.ModuleType("librosa")
        filt = types.ModuleType("librosa.filters")
        util = types.ModuleType("librosa.util")
        filt.mel = lambda *a, **k: None
        util.pad_center = lambda *a, **k: None
        util.tiny = lambda *a, **k: 1e-30
        lib.filters, lib.util = filt, util
        lib.stft = lib.istft = lambda *a, **k: None
        sys.modules.update({"librosa": lib, "librosa.filters": filt, "librosa.util": util})
    if "monotonic_align" not in sys.modules:
        # models.py:12 does `import monotonic_align`; give it a shim so the un-built
        # package under /root/reference is never imported (its __init__ needs the .so).
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from oracle import mas as _mas
        shim = types.ModuleType("monotonic_align")
        shim.maximum_path = lambda value, mask: _mas.maximum_path(value, mask, impl="ref")
        sys.modules["monotonic_align"] = shim
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        import commons, modules, attentions, models  # noqa: E401
    finally:
        sys.path.remove(REFERENCE_ROOT)
    _cached = types.SimpleNamespace(commons=commons, modules=modules, attentions=attentions, models=models)
    return _cached
