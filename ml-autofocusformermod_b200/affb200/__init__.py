"""affb200 — B200 (sm_100a) implementation of AutoFocusFormer's off-grid attention hot path.

Importing this package loads libaffb200.so and raises RuntimeError when it is missing (no CPU fallback),
like the reference's clusten/clusten.py:8-14.
"""
from . import _native  # noqa: F401  (loads the library; raises if it is missing)
from .clusten import CLUSTENQKFunction, CLUSTENAVFunction, CLUSTENWFFunction  # noqa: F401

__all__ = ["CLUSTENQKFunction", "CLUSTENAVFunction", "CLUSTENWFFunction"]
