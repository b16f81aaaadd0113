"""geopurify_b200 -- B200-native (sm_100a) implementation of GeoPurify's inference hot path.

Host code is Python/PyTorch and mirrors the reference's call surface; every stage runs in
libgeopurify_b200.so (hand-written CUDA behind a C ABI, include/geopurify_b200.h).  No CPU fallback.
"""
from . import _lib
from ._lib import GeoPurifyError

__all__ = ["_lib", "GeoPurifyError"]
