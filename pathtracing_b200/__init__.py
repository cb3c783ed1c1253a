"""pathtracing_b200 — B200-native (sm_100a CUDA) backend for chairclr/PathTracing's per-pixel
tracing path, behind a C ABI (include/ptb200.h, pathtracing_b200/libptb200.so).

Importing the package does not load the native library; constructing a tracer does, and fails
loudly if it is missing (there is no CPU fallback).
"""
from ._lib import LIB_PATH, PtbError, build, load   # noqa: F401
from .tracer import NativeTracer                    # noqa: F401
from .renderer import Renderer, Window              # noqa: F401

__version__ = "0.1.0"
