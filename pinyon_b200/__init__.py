"""pinyon_b200 — B200-native Monte-Carlo leaf evaluator for Pinyon (baskuit/pinyon).

The product is libpinyon_cuda.so (csrc/, C ABI in include/pinyon_cuda.h) and the
header-only C++ model in include/pinyon_cuda/.  This Python package only builds the
library and binds it for tests and benchmarks.
"""
from .build import build_library  # noqa: F401
from .capi import Context, Engine, PcuError, load_library  # noqa: F401

__all__ = ["build_library", "Context", "Engine", "PcuError", "load_library"]
