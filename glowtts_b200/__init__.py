"""glowtts_b200 -- B200-native (sm_100a) Glow-TTS hot path behind the reference's Python API.

    from glowtts_b200 import monotonic_align, commons, modules, attentions, models

The CUDA library (glowtts_b200/lib/libglowtts_b200.so, built by ``python -m glowtts_b200.build``)
is loaded on first use; there is no CPU fallback.
"""
__all__ = ["monotonic_align", "commons", "modules", "attentions", "models", "ops"]


def __getattr__(name):
    if name in __all__:
        import importlib
        return importlib.import_module("." + name, __name__)
    raise AttributeError(name)
