"""qwen3-asr.cpp_b200 — B200-native log-mel frontend + audio-encoder forward.

The product is the C-ABI shared library `libq3a.so` (include/q3a.h) built from csrc/ by
build.py; this package holds only that library's sources, the reference-compatible C++
shims (host/), a ctypes binding used by tests/bench (binding.py) and the synthetic-data
helpers (synth.py).  There is no CPU fallback: binding.load() raises if the library is missing.
"""
from . import synth  # noqa: F401
