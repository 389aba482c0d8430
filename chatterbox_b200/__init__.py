"""chatterbox_b200 — B200-native inference hot path of Chatterbox TTS behind the reference's own API.

    from chatterbox_b200 import ChatterboxTTS, ChatterboxVC

Submodules import torch lazily enough for CPU-only tooling (build, packers, tests); constructing any model
class requires a CUDA device and the in-tree `libcbx.so` — there is no CPU fallback.
"""

__all__ = ["ChatterboxTTS", "ChatterboxVC", "Conditionals", "punc_norm"]


def __getattr__(name):
    if name in ("ChatterboxTTS", "Conditionals", "punc_norm"):
        from . import tts
        return getattr(tts, name)
    if name == "ChatterboxVC":
        from .vc import ChatterboxVC
        return ChatterboxVC
    raise AttributeError(name)
