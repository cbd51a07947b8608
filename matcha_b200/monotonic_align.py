"""Drop-in for `matcha.utils.monotonic_align` (reference monotonic_align/__init__.py:7-22): `maximum_path(value, mask)`.

The reference moves the (B, Lx, T) log-prior to the host, runs its Cython/OpenMP dynamic programme and moves the path
back, every training step (matcha_tts.py:192).  Here the search runs where the tensors are, one CTA per utterance
(csrc/mas.cu), and returns the same 0/1 path bit for bit.  CUDA tensors only; there is no CPU fallback.
"""
from . import _lib


def maximum_path(value, mask):
    """value, mask: (B, t_x, t_y) -> path (B, t_x, t_y), same dtype and device as `value`."""
    return _lib.maximum_path(value, mask)
