"""ctypes wrappers of the CPU oracle (oracle/libhmj_oracle.so) and of the compiled reference
(oracle/_ref/libhmj_ref.so).  Test infrastructure: only tests/, smoke() and bench.py's CPU
legs import this module; the product (hashmergejoin_b200/) never does."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_PATH = os.path.join(ROOT, "oracle", "libhmj_oracle.so")
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libhmj_ref.so")

_P, _SZ, _I, _U64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64
_U64P = C.POINTER(C.c_uint64)


class OracleKeys(C.Structure):
    _fields_ = [("kind", C.c_int), ("u64", C.c_void_p), ("offsets", C.c_void_p), ("chars", C.c_void_p)]


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


class Keys:
    """Key storage of one relation (row-id indexed): u64 array or (offsets, chars)."""

    def __init__(self, u64=None, offsets=None, chars=None):
        self.u64 = _c(u64, np.uint64) if u64 is not None else None
        self.offsets = _c(offsets, np.uint64) if offsets is not None else None
        self.chars = _c(chars, np.uint8) if chars is not None else None
        self.kind = 0 if u64 is not None else 1
        self.n = self.u64.size if self.kind == 0 else self.offsets.size - 1

    def struct(self):
        return OracleKeys(self.kind, _p(self.u64), _p(self.offsets), _p(self.chars))


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        L = lib
        L.oracle_hash_bytes.restype = _U64
        L.oracle_hash_bytes.argtypes = [_P, _SZ, _U64]
        L.oracle_hash_strings.argtypes = [_P, _P, _SZ, _P]
        L.oracle_optimal_partition.restype = _I
        L.oracle_optimal_partition.argtypes = [_SZ]
        for f in (L.oracle_radix_sort_non_inplace, L.oracle_radix_sort_inplace_seq):
            f.restype = None
            f.argtypes = [_P, _P, _SZ, _I, C.POINTER(OracleKeys)]
        L.oracle_canonical_sort.argtypes = [_P, _P, _SZ]
        L.oracle_merge_walk_reference.restype = _SZ
        L.oracle_merge_walk_reference.argtypes = [_P, _P, _SZ, _P, _P, _SZ, C.POINTER(OracleKeys),
                                                  C.POINTER(OracleKeys), _P, _P, _SZ]
        L.oracle_equi_join.restype = _SZ
        L.oracle_equi_join.argtypes = [_P, _P, _SZ, _P, _P, _SZ, C.POINTER(OracleKeys),
                                       C.POINTER(OracleKeys), _P, _P, _SZ, _U64P]
        L.oracle_sort_pairs.argtypes = [_P, _P, _SZ]
        L.oracle_partition_counts.argtypes = [_P, _SZ, _I, _P]
        L.oracle_hash_join_raw.argtypes = [_P, _SZ, C.POINTER(OracleKeys), _P, _P, _SZ,
                                           C.POINTER(OracleKeys), _P, _U64P, _U64P]
        L.oracle_join_u64.restype = _SZ
        L.oracle_join_u64.argtypes = [_P, _SZ, _P, _SZ, _P, _P, _SZ, _U64P]

    def hash_bytes(self, b):
        buf = (C.c_char * max(len(b), 1)).from_buffer_copy(b if b else b"\0")
        return self.lib.oracle_hash_bytes(C.cast(buf, _P), len(b), 0xc70f6907)

    def hash_strings(self, offsets, chars):
        offsets, chars = _c(offsets, np.uint64), _c(chars, np.uint8)
        out = np.empty(offsets.size - 1, np.uint64)
        self.lib.oracle_hash_strings(_p(offsets), _p(chars), out.size, _p(out))
        return out

    def hashes(self, keys):
        return keys.u64.copy() if keys.kind == 0 else self.hash_strings(keys.offsets, keys.chars)

    def optimal_partition(self, n):
        return self.lib.oracle_optimal_partition(n)

    def radix_sort(self, hash_, keys, bits=0, rowid=None, inplace_seq=False):
        """reference output order of radix_non_inplace_par / radix_inplace_seq"""
        h = _c(hash_, np.uint64).copy()
        r = np.arange(h.size, dtype=np.uint32) if rowid is None else _c(rowid, np.uint32).copy()
        ks = keys.struct()
        f = self.lib.oracle_radix_sort_inplace_seq if inplace_seq else self.lib.oracle_radix_sort_non_inplace
        f(_p(h), _p(r), h.size, bits, C.byref(ks))
        return h, r

    def canonical(self, hash_, rowid):
        h, r = _c(hash_, np.uint64).copy(), _c(rowid, np.uint32).copy()
        self.lib.oracle_canonical_sort(_p(h), _p(r), h.size)
        return h, r

    def _join(self, fn, rh, rr, sh, sr, rkeys, skeys, with_sum):
        rh, sh = _c(rh, np.uint64), _c(sh, np.uint64)
        rr, sr = _c(rr, np.uint32), _c(sr, np.uint32)
        rk, sk = rkeys.struct(), skeys.struct()
        args = [_p(rh), _p(rr), rh.size, _p(sh), _p(sr), sh.size, C.byref(rk), C.byref(sk)]
        s = C.c_uint64()
        extra = [C.byref(s)] if with_sum else []
        m = fn(*args, None, None, 0, *extra)
        out_r, out_s = np.empty(m, np.uint32), np.empty(m, np.uint32)
        m2 = fn(*args, _p(out_r), _p(out_s), m, *extra)
        assert m2 == m
        return out_r, out_s, s.value

    def merge_walk_reference(self, rh, rr, sh, sr, rkeys, skeys):
        r, s, _ = self._join(self.lib.oracle_merge_walk_reference, rh, rr, sh, sr, rkeys, skeys, False)
        return r, s

    def equi_join(self, rh, rr, sh, sr, rkeys, skeys):
        return self._join(self.lib.oracle_equi_join, rh, rr, sh, sr, rkeys, skeys, True)

    def sort_pairs(self, out_r, out_s):
        a, b = _c(out_r, np.uint32).copy(), _c(out_s, np.uint32).copy()
        self.lib.oracle_sort_pairs(_p(a), _p(b), a.size)
        return a, b

    def partition_counts(self, hash_, bits):
        h = _c(hash_, np.uint64)
        out = np.zeros(1 << bits, np.uint64)
        self.lib.oracle_partition_counts(_p(h), h.size, bits, _p(out))
        return out

    def hash_join_raw(self, rkeys, skeys, rvals=None, svals=None):
        rh, sh = self.hashes(rkeys), self.hashes(skeys)
        rk, sk = rkeys.struct(), skeys.struct()
        hits, sm = C.c_uint64(), C.c_uint64()
        rv = _c(rvals, np.uint64) if rvals is not None else None
        sv = _c(svals, np.uint64) if svals is not None else None
        self.lib.oracle_hash_join_raw(_p(rh), rh.size, C.byref(rk), _p(rv), _p(sh), sh.size,
                                      C.byref(sk), _p(sv), C.byref(hits), C.byref(sm))
        return hits.value, sm.value

    def full_join(self, rkeys, skeys):
        """hash + reference-order sort + TRUE equi-join; canonical (sorted) pairs and the sum."""
        rh, rr = self.radix_sort(self.hashes(rkeys), rkeys)
        sh, sr = self.radix_sort(self.hashes(skeys), skeys)
        out_r, out_s, sm = self.equi_join(rh, rr, sh, sr, rkeys, skeys)
        out_r, out_s = self.sort_pairs(out_r, out_s)
        return out_r, out_s, sm


class Ref:
    """The compiled, unmodified reference (oracle/_ref/libhmj_ref.so)."""

    def __init__(self, lib):
        self.lib = lib
        L = lib
        L.ref_hardware_concurrency.restype = _I
        L.ref_optimal_partition.restype = _I
        L.ref_optimal_partition.argtypes = [_SZ]
        L.ref_std_hash_bytes.restype = _U64
        L.ref_std_hash_bytes.argtypes = [_P, _SZ]
        L.ref_std_hash_u64.restype = _U64
        L.ref_std_hash_u64.argtypes = [_U64]
        L.ref_hash_strings.argtypes = [_P, _P, _SZ, _P]
        L.ref_sort_u64.restype = C.c_double
        L.ref_sort_u64.argtypes = [_P, _SZ, _I, _I, _P, _P]
        L.ref_sort_strings.restype = C.c_double
        L.ref_sort_strings.argtypes = [_P, _P, _SZ, _I, _I, _P, _P]
        L.ref_sort_inplace.restype = C.c_double
        L.ref_sort_inplace.argtypes = [_P, _P, _SZ, _I, _I, _I]
        L.ref_sort_int_pairs.restype = C.c_double
        L.ref_sort_int_pairs.argtypes = [_P, _P, _SZ, _I, _I]
        L.ref_join_u64.restype = C.c_double
        L.ref_join_u64.argtypes = [_P, _SZ, _P, _SZ, _I, _P, _P, _SZ, _U64P, _U64P]
        L.ref_join_strings.restype = C.c_double
        L.ref_join_strings.argtypes = [_P, _P, _SZ, _P, _P, _SZ, _I, _P, _P, _SZ, _U64P, _U64P]
        L.ref_hash_join_raw_u64.restype = C.c_double
        L.ref_hash_join_raw_u64.argtypes = [_P, _SZ, _P, _SZ, _U64P]
        L.ref_hash_join_raw_strings.restype = C.c_double
        L.ref_hash_join_raw_strings.argtypes = [_P, _P, _SZ, _P, _P, _SZ, _U64P]
        L.ref_partition_counts_u64.argtypes = [_P, _SZ, _I, _I, _P]
        L.ref_partition_counts_strings.argtypes = [_P, _P, _SZ, _I, _I, _P]

    def cores(self):
        return self.lib.ref_hardware_concurrency()

    def hash_bytes(self, b):
        buf = (C.c_char * max(len(b), 1)).from_buffer_copy(b if b else b"\0")
        return self.lib.ref_std_hash_bytes(C.cast(buf, _P), len(b))

    def hash_strings(self, offsets, chars):
        offsets, chars = _c(offsets, np.uint64), _c(chars, np.uint8)
        out = np.empty(offsets.size - 1, np.uint64)
        self.lib.ref_hash_strings(_p(offsets), _p(chars), out.size, _p(out))
        return out

    def sort(self, keys, threads=1, bits=0):
        """radix_non_inplace_par -> (hash, rowid, seconds)"""
        h, r = np.empty(keys.n, np.uint64), np.empty(keys.n, np.uint32)
        if keys.kind == 0:
            t = self.lib.ref_sort_u64(_p(keys.u64), keys.n, threads, bits, _p(h), _p(r))
        else:
            t = self.lib.ref_sort_strings(_p(keys.offsets), _p(keys.chars), keys.n, threads, bits, _p(h), _p(r))
        return h, r, t

    def sort_inplace(self, hash_, mode, threads=1, bits=0):
        h = _c(hash_, np.uint64).copy()
        r = np.arange(h.size, dtype=np.uint32)
        t = self.lib.ref_sort_inplace(_p(h), _p(r), h.size, mode, threads, bits)
        return h, r, t

    def sort_int_pairs(self, keys, vals, mode, threads=1):
        k, v = _c(keys, np.uint64).copy(), _c(vals, np.uint32).copy()
        t = self.lib.ref_sort_int_pairs(_p(k), _p(v), k.size, mode, threads)
        return k, v, t

    def join(self, rkeys, skeys, threads=1, want_pairs=True):
        """HashMergeJoin + full iteration -> (out_r, out_s, sum, seconds)"""
        m, sm = C.c_uint64(), C.c_uint64()
        cap = min(rkeys.n, skeys.n) * 4 + 16 if want_pairs else 0
        out_r = np.empty(cap, np.uint32) if want_pairs else None
        out_s = np.empty(cap, np.uint32) if want_pairs else None
        if rkeys.kind == 0:
            t = self.lib.ref_join_u64(_p(rkeys.u64), rkeys.n, _p(skeys.u64), skeys.n, threads,
                                      _p(out_r), _p(out_s), cap, C.byref(m), C.byref(sm))
        else:
            t = self.lib.ref_join_strings(_p(rkeys.offsets), _p(rkeys.chars), rkeys.n,
                                          _p(skeys.offsets), _p(skeys.chars), skeys.n, threads,
                                          _p(out_r), _p(out_s), cap, C.byref(m), C.byref(sm))
        if want_pairs:
            assert m.value <= cap
            return out_r[: m.value], out_s[: m.value], sm.value, t
        return m.value, None, sm.value, t

    def hash_join_raw(self, rkeys, skeys):
        """unordered_map build over S + probe with R (hashjoin_bench.cc:41-53) -> (sum, seconds)"""
        sm = C.c_uint64()
        if rkeys.kind == 0:
            t = self.lib.ref_hash_join_raw_u64(_p(rkeys.u64), rkeys.n, _p(skeys.u64), skeys.n, C.byref(sm))
        else:
            t = self.lib.ref_hash_join_raw_strings(_p(rkeys.offsets), _p(rkeys.chars), rkeys.n,
                                                   _p(skeys.offsets), _p(skeys.chars), skeys.n, C.byref(sm))
        return sm.value, t

    def partition_counts(self, keys, threads, bits):
        out = np.zeros(1 << bits, np.uint64)
        if keys.kind == 0:
            self.lib.ref_partition_counts_u64(_p(keys.u64), keys.n, threads, bits, _p(out))
        else:
            self.lib.ref_partition_counts_strings(_p(keys.offsets), _p(keys.chars), keys.n, threads, bits, _p(out))
        return out


def load_oracle():
    if not os.path.exists(ORACLE_PATH):
        raise RuntimeError(f"{ORACLE_PATH} missing: run `make oracle`")
    return Oracle(C.CDLL(ORACLE_PATH))


def load_ref():
    if not os.path.exists(REF_PATH):
        return None
    return Ref(C.CDLL(REF_PATH))
