"""ctypes binding of libhmj.so (the C ABI declared in include/hmj.h).

There is no fallback of any kind: if the CUDA library has not been built (``make lib`` or
``__graft_entry__.build()``) importing the binding raises, and every call that returns a
non-zero status raises :class:`HmjError` with the library's own message.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# HMJ_LIB_PATH: an instrumented build of the same library (scripts/trace_pass.py), never a fallback
LIB_PATH = os.environ.get("HMJ_LIB_PATH") or os.path.join(_HERE, "libhmj.so")

HMJ_OK = 0
HMJ_EARG, HMJ_ECUDA, HMJ_ENOMEM, HMJ_ENODEVICE, HMJ_ECAPACITY = -1, -2, -3, -4, -5
KEY_U64, KEY_STRING = 0, 1

_STATUS = {
    HMJ_EARG: "HMJ_EARG", HMJ_ECUDA: "HMJ_ECUDA", HMJ_ENOMEM: "HMJ_ENOMEM",
    HMJ_ENODEVICE: "HMJ_ENODEVICE", HMJ_ECAPACITY: "HMJ_ECAPACITY",
}


class HmjError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{_STATUS.get(code, code)}: {msg}")
        self.code = code


class Relation(C.Structure):
    """struct hmj_relation"""
    _fields_ = [
        ("d_hash", C.c_void_p), ("d_rowid", C.c_void_p), ("n", C.c_size_t),
        ("key_kind", C.c_int), ("d_offsets", C.c_void_p), ("d_chars", C.c_void_p),
    ]


class Timings(C.Structure):
    """struct hmj_timings"""
    _fields_ = [
        ("hash_ms", C.c_float), ("histogram_ms", C.c_float), ("sort_ms", C.c_float),
        ("join_ms", C.c_float), ("h2d_ms", C.c_float), ("d2h_ms", C.c_float),
        ("total_ms", C.c_float), ("sort_passes", C.c_uint32),
        ("pass_ms_sum", C.c_float), ("pass_launches", C.c_uint32),
        ("sort_finishes", C.c_uint32), ("sort_fallbacks", C.c_uint32),
        ("sort_uniform", C.c_uint32), ("kernel_launches", C.c_uint32 * 8),
        ("kernel_ms", C.c_float * 8), ("kernel_bytes", C.c_double * 8),
        ("bucket_join", C.c_uint32), ("scatter_levels", C.c_uint32),
    ]


_P, _SZ, _I = C.c_void_p, C.c_size_t, C.c_int
_U64P = C.POINTER(C.c_uint64)
_RELP = C.POINTER(Relation)

# name -> (restype, argtypes); must list every function include/hmj.h declares
SIGNATURES = {
    "hmj_create": (_I, [C.POINTER(_P), _I]),
    "hmj_create_on_stream": (_I, [C.POINTER(_P), _I, _P]),
    "hmj_destroy": (_I, [_P]),
    "hmj_last_error": (C.c_char_p, [_P]),
    "hmj_synchronize": (_I, [_P]),
    "hmj_version": (C.c_char_p, []),
    "hmj_launch_count": (_I, [_P, _U64P]),
    "hmj_release_workspace": (_I, [_P]),
    "hmj_optimal_partition": (_I, [_SZ]),
    "hmj_hash_u64": (_I, [_P, _P, _SZ, _P]),
    "hmj_hash_strings": (_I, [_P, _P, _P, _SZ, _P]),
    "hmj_sort_pairs": (_I, [_P, _P, _P, _SZ, _P, _P]),
    "hmj_hash_sort_u64": (_I, [_P, _P, _SZ, _P, _P]),
    "hmj_hash_sort_strings": (_I, [_P, _P, _P, _SZ, _P, _P]),
    "hmj_sort_int_pairs": (_I, [_P, _P, _P, _SZ, _P, _P]),
    "hmj_merge_join_count": (_I, [_P, _RELP, _RELP, _U64P]),
    "hmj_merge_join_write": (_I, [_P, _RELP, _RELP, _P, _P, _SZ, _U64P]),
    "hmj_merge_join_aggregate": (_I, [_P, _RELP, _RELP, _P, _P, _U64P, _U64P]),
    "hmj_join_u64": (_I, [_P, _P, _SZ, _P, _SZ, _U64P]),
    "hmj_join_pairs": (_I, [_P, _P, _P, _SZ, _P, _P, _SZ, _I, _U64P]),
    "hmj_join_strings": (_I, [_P, _P, _P, _SZ, _P, _P, _SZ, _U64P]),
    "hmj_join_strings_range": (_I, [_P, _P, _P, _SZ, _P, _P, _SZ, _I, _U64P]),
    "hmj_gather_strings": (_I, [_P, _P, _P, _P, _SZ, _P, _P]),
    "hmj_string_row_spans": (_I, [_P, _P, _P, _SZ, _P, _P]),
    "hmj_gather_spans": (_I, [_P, _P, _P, _SZ, _P, _P]),
    "hmj_set_join_order": (_I, [_P, _I]),
    "hmj_join_result": (_I, [_P, C.POINTER(_P), C.POINTER(_P), _U64P]),
    "hmj_join_result_copy": (_I, [_P, _P, _P, _SZ]),
    "hmj_join_result_to_host": (_I, [_P, _P, _P, _SZ]),
    "hmj_join_u64_host": (_I, [_P, _P, _SZ, _P, _SZ, _P, _P, _SZ, _U64P]),
    "hmj_join_strings_host": (_I, [_P, _P, _P, _SZ, _P, _P, _SZ, _P, _P, _SZ, _U64P]),
    "hmj_host_alloc": (_I, [C.POINTER(_P), _SZ]),
    "hmj_host_free": (_I, [_P]),
    "hmj_hash_sort_u64_host": (_I, [_P, _P, _SZ, _P, _P]),
    "hmj_hash_sort_strings_host": (_I, [_P, _P, _P, _SZ, _P, _P]),
    "hmj_sort_pairs_host": (_I, [_P, _P, _P, _SZ, _P, _P]),
    "hmj_partition_counts": (_I, [_P, _P, _SZ, _I, _P]),
    "hmj_partition_pairs": (_I, [_P, _P, _P, _SZ, _I, _P, _P, _P]),
    "hmj_partition_pairs_base": (_I, [_P, _P, _P, C.c_uint32, _SZ, _I, _P, _P, _P]),
    "hmj_partition_pairs_noalloc": (_I, [_P, _P, _P, C.c_uint32, _SZ, _I, _P, _P, _P]),
    "hmj_debug_plan_from_sample": (_I, [_P, C.c_uint32, _SZ, _SZ, _P]),
    "hmj_set_profiling": (_I, [_P, _I]),
    "hmj_partition_counts_host": (_I, [_P, _P, _SZ, _I, _P]),
    "hmj_hashjoin_build_probe": (_I, [_P, _RELP, _RELP, _P, _P, _U64P, _U64P]),
    "hmj_get_timings": (_I, [_P, C.POINTER(Timings)]),
    "hmj_dist_setup": (_I, [_P, _I, _I, _SZ, _P]),
    "hmj_dist_connect": (_I, [_P, _P]),
    "hmj_dist_handle_bytes": (_I, []),
    "hmj_dist_join_u64": (_I, [_P, _P, _SZ, C.c_uint32, _P, _SZ, C.c_uint32, _U64P, C.POINTER(C.c_int)]),
    "hmj_dist_last_exchange": (_I, [_P, _U64P, _U64P, _U64P]),
    "hmj_dist_teardown": (_I, [_P]),
    "hmj_xchg_handle_bytes": (_I, []),
    "hmj_xchg_setup": (_I, [_P, _I, _I, _SZ, _SZ, _P]),
    "hmj_xchg_connect": (_I, [_P, _P]),
    "hmj_xchg_send": (_I, [_P, _I, _P, _P, _P, _P, _P, _SZ, _P, _P, _P]),
    "hmj_xchg_buffers": (_I, [_P, _I, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), C.POINTER(_SZ), C.POINTER(_SZ)]),
    "hmj_xchg_teardown": (_I, [_P]),
    "hmj_gen_permute": (_I, [_P, _P, C.c_uint32, _SZ, C.c_uint32, C.c_uint32, _P]),
    "hmj_gen_zipf": (_I, [_P, _P, C.c_uint32, C.c_uint64, C.c_uint64, _SZ, _P]),
    "hmj_gen_strgen_lengths": (_I, [_P, _P, C.c_uint32, _P, _SZ, _P]),
    "hmj_gen_strgen_write": (_I, [_P, _P, _P, C.c_uint32, _P, _SZ, _P, _P]),
}

_lib = None


def load():
    """Loads libhmj.so (once). Raises ImportError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make lib` (or __graft_entry__.build()). "
            "hashmergejoin_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
