"""matcha_b200: B200-native (sm_100a) synthesis hot path of Matcha-TTS -- CFM Euler decode + HiFi-GAN v1 vocoder.

Public surface mirrors the reference's (Mocifiel/Matcha): `CFM`, `Decoder`, `Generator`, `MatchaTTS.synthesise`.
The compute lives in libmatcha_b200.so (hand-written CUDA, C ABI in include/matcha_b200.h); importing this package
does not need a GPU, calling into it does.
"""
from ._lib import Plan, euler_schedule  # noqa: F401
from .cfm import BASECFM, CFM, Decoder  # noqa: F401
from .hifigan import AttrDict, Denoiser, Generator, V1 as HIFIGAN_V1  # noqa: F401
from .matcha_tts import MatchaTTS, TextEncoder  # noqa: F401
from . import monotonic_align  # noqa: F401

__all__ = ["Plan", "euler_schedule", "BASECFM", "CFM", "Decoder", "Generator", "Denoiser", "AttrDict", "HIFIGAN_V1", "MatchaTTS",
           "TextEncoder", "monotonic_align"]
