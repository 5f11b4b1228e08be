"""hashmergejoin_b200 -- B200 (sm_100a) hash -> radix-sort -> merge-join hot path of
dryman/HashMergeJoin behind a C ABI (include/hmj.h, hashmergejoin_b200/libhmj.so).

Layout: csrc/ CUDA kernels + C ABI; facade/ C++ header templates keeping the reference's names
(radix_hash.h, hashjoin.h, partitioned_hash.h, radix_sort.h); api.py ctypes plumbing used by the
tests and bench.py; datagen.py seeded synthetic relations.  There is no CPU fallback.
"""
from ._lib import KEY_STRING, KEY_U64, HmjError, LIB_PATH, load  # noqa: F401
from .api import Context, SortedRelation, optimal_partition  # noqa: F401

__all__ = ["Context", "SortedRelation", "HmjError", "KEY_U64", "KEY_STRING", "LIB_PATH", "load",
           "optimal_partition"]
