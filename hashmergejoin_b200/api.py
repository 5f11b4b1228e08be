"""Host-side Python plumbing over the C ABI: device memory comes from torch tensors, host
buffers from numpy arrays.  Nothing here computes: every method is one C-ABI call into the CUDA
library (hashmergejoin_b200/csrc).  Columns are u64/u32; torch has no arithmetic on unsigned
64-bit, so device columns are int64/int32 tensors that are reinterpreted, never converted.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KEY_STRING, KEY_U64, HmjError, Relation, Timings


def _dptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _hptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def _check_dev(t, dtype_bytes):
    assert t.is_cuda and t.is_contiguous() and t.element_size() == dtype_bytes, \
        "expected a contiguous CUDA tensor of %d-byte elements" % dtype_bytes


class _DeviceSpan:
    """A raw device allocation of the library seen through __cuda_array_interface__, so that
    torch.as_tensor can wrap it without a copy."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class SortedRelation:
    """Sorted (hash,row-id) columns of one relation plus the key storage the row ids point to."""

    def __init__(self, hash_, rowid, kind=KEY_U64, offsets=None, chars=None):
        self.hash, self.rowid, self.kind = hash_, rowid, kind
        self.offsets, self.chars = offsets, chars

    def as_struct(self):
        r = Relation()
        r.d_hash = self.hash.data_ptr()
        r.d_rowid = self.rowid.data_ptr() if self.rowid is not None else None
        r.n = self.hash.numel()
        r.key_kind = self.kind
        r.d_offsets = self.offsets.data_ptr() if self.offsets is not None else None
        r.d_chars = self.chars.data_ptr() if self.chars is not None else None
        return r


class Context:
    """One device + one stream (hmj_ctx).  With ``use_torch_stream`` work is issued on torch's
    current stream so that torch.cuda.Event timing and torch allocations are ordered with it."""

    def __init__(self, device=0, use_torch_stream=True):
        import torch
        self._torch = torch
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise HmjError(_lib.HMJ_ENODEVICE, "no CUDA device: hashmergejoin_b200 has no CPU path")
        self.device = torch.device("cuda", device)
        h = C.c_void_p()
        if use_torch_stream:
            with torch.cuda.device(self.device):
                stream = torch.cuda.current_stream().cuda_stream
            rc = self.lib.hmj_create_on_stream(C.byref(h), device, C.c_void_p(stream))
        else:
            rc = self.lib.hmj_create(C.byref(h), device)
        if rc != 0:
            raise HmjError(rc, "hmj_create failed")
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.hmj_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise HmjError(rc, self.lib.hmj_last_error(self.h).decode())

    def _empty(self, n, dtype):
        return self._torch.empty(int(n), dtype=dtype, device=self.device)

    # ---- A1 hashing
    def hash_u64(self, keys):
        _check_dev(keys, 8)
        out = self._empty(keys.numel(), self._torch.int64)
        self._ck(self.lib.hmj_hash_u64(self.h, _dptr(keys), keys.numel(), _dptr(out)))
        return out

    def hash_strings(self, offsets, chars):
        _check_dev(offsets, 8)
        n = offsets.numel() - 1
        out = self._empty(n, self._torch.int64)
        self._ck(self.lib.hmj_hash_strings(self.h, _dptr(offsets), _dptr(chars), n, _dptr(out)))
        return out

    # ---- A2-A4 sorting
    def sort_pairs(self, hash_, rowid=None, out_hash=None, out_rowid=None):
        _check_dev(hash_, 8)
        n = hash_.numel()
        out_hash = out_hash if out_hash is not None else self._empty(n, self._torch.int64)
        out_rowid = out_rowid if out_rowid is not None else self._empty(n, self._torch.int32)
        self._ck(self.lib.hmj_sort_pairs(self.h, _dptr(hash_), _dptr(rowid), n, _dptr(out_hash),
                                         _dptr(out_rowid)))
        return out_hash, out_rowid

    def hash_sort_u64(self, keys, out_hash=None, out_rowid=None):
        _check_dev(keys, 8)
        n = keys.numel()
        out_hash = out_hash if out_hash is not None else self._empty(n, self._torch.int64)
        out_rowid = out_rowid if out_rowid is not None else self._empty(n, self._torch.int32)
        self._ck(self.lib.hmj_hash_sort_u64(self.h, _dptr(keys), n, _dptr(out_hash), _dptr(out_rowid)))
        return out_hash, out_rowid

    def hash_sort_strings(self, offsets, chars, out_hash=None, out_rowid=None):
        n = offsets.numel() - 1
        out_hash = out_hash if out_hash is not None else self._empty(n, self._torch.int64)
        out_rowid = out_rowid if out_rowid is not None else self._empty(n, self._torch.int32)
        self._ck(self.lib.hmj_hash_sort_strings(self.h, _dptr(offsets), _dptr(chars), n,
                                                _dptr(out_hash), _dptr(out_rowid)))
        return out_hash, out_rowid

    # ---- A5 merge join on sorted relations
    def merge_join_count(self, R, S):
        m = C.c_uint64()
        r, s = R.as_struct(), S.as_struct()
        self._ck(self.lib.hmj_merge_join_count(self.h, C.byref(r), C.byref(s), C.byref(m)))
        return m.value

    def merge_join(self, R, S, cap=None):
        """Fused single-pass join.  First try with room for max(|R|,|S|) pairs (enough for any
        join with unique keys on one side); an overflowing join reports the exact count and is
        re-run into a buffer of that size."""
        nr, ns = R.hash.numel(), S.hash.numel()
        cap = max(nr, ns, 1) if cap is None else cap
        r, s = R.as_struct(), S.as_struct()
        while True:
            out_r = self._empty(cap, self._torch.int32)
            out_s = self._empty(cap, self._torch.int32)
            m = C.c_uint64()
            rc = self.lib.hmj_merge_join_write(self.h, C.byref(r), C.byref(s), _dptr(out_r),
                                               _dptr(out_s), cap, C.byref(m))
            if rc == _lib.HMJ_ECAPACITY and m.value > cap:
                cap = m.value
                continue
            self._ck(rc)
            return out_r[:m.value], out_s[:m.value]

    def merge_join_aggregate(self, R, S, rvals=None, svals=None):
        m, sm = C.c_uint64(), C.c_uint64()
        r, s = R.as_struct(), S.as_struct()
        self._ck(self.lib.hmj_merge_join_aggregate(self.h, C.byref(r), C.byref(s), _dptr(rvals),
                                                   _dptr(svals), C.byref(m), C.byref(sm)))
        return m.value, sm.value

    # ---- whole operator, device inputs
    def join_u64(self, rkeys, skeys):
        _check_dev(rkeys, 8)
        _check_dev(skeys, 8)
        m = C.c_uint64()
        self._ck(self.lib.hmj_join_u64(self.h, _dptr(rkeys), rkeys.numel(), _dptr(skeys),
                                       skeys.numel(), C.byref(m)))
        return m.value

    def join_pairs(self, rhash, rrowid, shash, srowid, prefix_bits=0):
        """hmj_join_pairs: unsorted (hash, row id) relations of one hash range (multi-GPU local step)"""
        m = C.c_uint64()
        self._ck(self.lib.hmj_join_pairs(self.h, _dptr(rhash), _dptr(rrowid), rhash.numel(),
                                         _dptr(shash), _dptr(srowid), shash.numel(), prefix_bits,
                                         C.byref(m)))
        return m.value

    def set_join_order(self, hash_order):
        """True: pairs in ascending hash order, reproducible (sort + merge-path plan)"""
        self._ck(self.lib.hmj_set_join_order(self.h, 1 if hash_order else 0))

    def join_result(self):
        """device views (int32 tensors) of the last whole-operator result"""
        pr, ps, m = C.c_void_p(), C.c_void_p(), C.c_uint64()
        self._ck(self.lib.hmj_join_result(self.h, C.byref(pr), C.byref(ps), C.byref(m)))
        return pr.value, ps.value, m.value

    def join_result_tensors(self, m):
        """the last whole-operator result as two int32 CUDA tensors (hmj_join_result_copy)"""
        out_r, out_s = self._empty(m, self._torch.int32), self._empty(m, self._torch.int32)
        self._ck(self.lib.hmj_join_result_copy(self.h, _dptr(out_r), _dptr(out_s), m))
        return out_r, out_s

    # ---- multi-GPU join, exchange fused into the first scatter level (dist_join.cu)
    def dist_setup(self, rank, world, rows_total):
        handles = np.zeros(self.lib.hmj_dist_handle_bytes(), np.uint8)
        self._ck(self.lib.hmj_dist_setup(self.h, rank, world, rows_total, _hptr(handles)))
        return handles

    def dist_connect(self, all_handles):
        all_handles = np.ascontiguousarray(all_handles, np.uint8)
        self._ck(self.lib.hmj_dist_connect(self.h, _hptr(all_handles)))

    def dist_join_u64(self, rkeys, r_rowid_base, skeys, s_rowid_base):
        """hmj_dist_join_u64 -> (pairs of this rank's hash range, ok)"""
        _check_dev(rkeys, 8)
        _check_dev(skeys, 8)
        m, ok = C.c_uint64(), C.c_int()
        self._ck(self.lib.hmj_dist_join_u64(self.h, _dptr(rkeys), rkeys.numel(), r_rowid_base,
                                            _dptr(skeys), skeys.numel(), s_rowid_base,
                                            C.byref(m), C.byref(ok)))
        return m.value, ok.value

    def dist_last_exchange(self):
        """(pairs written into other ranks' memory, R pairs received, S pairs received)"""
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._ck(self.lib.hmj_dist_last_exchange(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def dist_teardown(self):
        self._ck(self.lib.hmj_dist_teardown(self.h))

    def join_strings(self, roff, rchars, soff, schars, prefix_bits=0):
        """hmj_join_strings[_range]: prefix_bits > 0 when every key hashes into one hash range
        (the local step of the multi-GPU join)"""
        m = C.c_uint64()
        self._ck(self.lib.hmj_join_strings_range(self.h, _dptr(roff), _dptr(rchars), roff.numel() - 1,
                                                 _dptr(soff), _dptr(schars), soff.numel() - 1,
                                                 prefix_bits, C.byref(m)))
        return m.value

    def gather_strings(self, offsets, chars, rows, want_lens=False):
        """keys of `rows` (int32 tensor of row ids) as a new (offsets, chars) column pair; with
        want_lens also the rows' lengths (int32) in that order"""
        t = self._torch
        _check_dev(offsets, 8)
        n = rows.numel()
        start = self._empty(n, t.int64)
        lens = self._empty(n, t.int32)
        self._ck(self.lib.hmj_string_row_spans(self.h, _dptr(offsets), _dptr(rows), n, _dptr(start), _dptr(lens)))
        out_off = t.zeros(n + 1, dtype=t.int64, device=offsets.device)
        if n:
            t.cumsum(lens, 0, dtype=t.int64, out=out_off[1:])
        total = int(out_off[-1].item()) if n else 0
        out_chars = self._empty(total + 16, t.uint8)[:total]
        self._ck(self.lib.hmj_gather_spans(self.h, _dptr(chars), _dptr(start), n, _dptr(out_off), _dptr(out_chars)))
        return (out_off, out_chars, lens) if want_lens else (out_off, out_chars)

    def join_result_to_host(self, m):
        out_r = np.empty(m, dtype=np.uint32)
        out_s = np.empty(m, dtype=np.uint32)
        self._ck(self.lib.hmj_join_result_to_host(self.h, _hptr(out_r), _hptr(out_s), m))
        return out_r, out_s

    # ---- whole operator, host inputs (numpy): the facade / e2e path
    def join_u64_host(self, rkeys, skeys, out_r=None, out_s=None):
        m = C.c_uint64()
        cap = 0 if out_r is None else out_r.size
        self._ck(self.lib.hmj_join_u64_host(self.h, _hptr(rkeys), rkeys.size, _hptr(skeys), skeys.size,
                                            _hptr(out_r), _hptr(out_s), cap, C.byref(m)))
        return m.value

    def join_strings_host(self, roff, rchars, soff, schars, out_r=None, out_s=None):
        m = C.c_uint64()
        cap = 0 if out_r is None else out_r.size
        self._ck(self.lib.hmj_join_strings_host(self.h, _hptr(roff), _hptr(rchars), roff.size - 1,
                                                _hptr(soff), _hptr(schars), soff.size - 1,
                                                _hptr(out_r), _hptr(out_s), cap, C.byref(m)))
        return m.value

    def hash_sort_u64_host(self, keys, out_hash=None, out_rowid=None):
        n = keys.size
        h = out_hash if out_hash is not None else np.empty(n, np.uint64)
        r = out_rowid if out_rowid is not None else np.empty(n, np.uint32)
        self._ck(self.lib.hmj_hash_sort_u64_host(self.h, _hptr(keys), n, _hptr(h), _hptr(r)))
        return h, r

    def hash_sort_strings_host(self, offsets, chars, out_hash=None, out_rowid=None):
        """out_hash / out_rowid: caller's (pinned) result arrays; fresh pageable ones otherwise"""
        n = offsets.size - 1
        h = out_hash if out_hash is not None else np.empty(n, np.uint64)
        r = out_rowid if out_rowid is not None else np.empty(n, np.uint32)
        self._ck(self.lib.hmj_hash_sort_strings_host(self.h, _hptr(offsets), _hptr(chars), n,
                                                     _hptr(h), _hptr(r)))
        return h, r

    def sort_pairs_host(self, hash_, rowid=None):
        n = hash_.size
        h, r = np.empty(n, np.uint64), np.empty(n, np.uint32)
        self._ck(self.lib.hmj_sort_pairs_host(self.h, _hptr(hash_), _hptr(rowid), n, _hptr(h), _hptr(r)))
        return h, r

    # ---- A6 partitioning
    def partition_counts_host(self, hash_, bits):
        counts = np.zeros(1 << bits, np.uint64)
        self._ck(self.lib.hmj_partition_counts_host(self.h, _hptr(hash_), hash_.size, bits, _hptr(counts)))
        return counts

    def partition_pairs(self, hash_, rowid, bits, rowid_base=0, out=None):
        n = hash_.numel()
        if out is None:
            out = (self._empty(n, self._torch.int64), self._empty(n, self._torch.int32),
                   self._empty((1 << bits) + 1, self._torch.int64))
        oh, orow, offs = out
        self._ck(self.lib.hmj_partition_pairs_base(self.h, _dptr(hash_), _dptr(rowid), rowid_base, n,
                                                   bits, _dptr(oh), _dptr(orow), _dptr(offs)))
        return oh, orow, offs

    def partition_pairs_noalloc(self, hash_, rowid, bits, out, rowid_base=0):
        """partition_pairs for callers that may not allocate device memory: the scratch is a fixed
        array inside the library (include/hmj.h, hmj_partition_pairs_noalloc), `out` =
        (hash_out[n] i64, rowid_out[n] i32, offsets[(1<<bits)+1] i64) is the caller's, and at most
        NOALLOC_MAX_ROWS rows are taken.  Nothing is allocated here either."""
        oh, orow, offs = out
        self._ck(self.lib.hmj_partition_pairs_noalloc(self.h, _dptr(hash_), _dptr(rowid), rowid_base,
                                                      hash_.numel(), bits, _dptr(oh), _dptr(orow),
                                                      _dptr(offs)))
        return oh, orow, offs

    NOALLOC_MAX_ROWS = 1 << 24

    # ---- A7 hash-join baseline
    def hashjoin_build_probe(self, R, S, rvals=None, svals=None):
        hits, sm = C.c_uint64(), C.c_uint64()
        r, s = R.as_struct(), S.as_struct()
        self._ck(self.lib.hmj_hashjoin_build_probe(self.h, C.byref(r), C.byref(s), _dptr(rvals),
                                                   _dptr(svals), C.byref(hits), C.byref(sm)))
        return hits.value, sm.value

    # ---- string keys over several GPUs: key gather fused with the exchange (dist_strings.cu)
    def string_row_spans(self, offsets, rows):
        """(start int64, lens int32) of the rows `rows` (int32 tensor) of a string column"""
        t = self._torch
        n = rows.numel()
        start, lens = self._empty(n, t.int64), self._empty(n, t.int32)
        self._ck(self.lib.hmj_string_row_spans(self.h, _dptr(offsets), _dptr(rows), n, _dptr(start), _dptr(lens)))
        return start, lens

    def xchg_setup(self, rank, world, rows_cap, bytes_cap):
        handles = np.zeros(self.lib.hmj_xchg_handle_bytes(), np.uint8)
        self._ck(self.lib.hmj_xchg_setup(self.h, rank, world, rows_cap, bytes_cap, _hptr(handles)))
        return handles

    def xchg_connect(self, all_handles):
        all_handles = np.ascontiguousarray(all_handles, np.uint8)
        self._ck(self.lib.hmj_xchg_connect(self.h, _hptr(all_handles)))

    def xchg_send(self, side, chars, start, send_off, rowid, lens, part_first, row_base, byte_base):
        a = [np.ascontiguousarray(x, np.uint64) for x in (part_first, row_base, byte_base)]
        self._ck(self.lib.hmj_xchg_send(self.h, side, _dptr(chars), _dptr(start), _dptr(send_off), _dptr(rowid),
                                        _dptr(lens), rowid.numel(), _hptr(a[0]), _hptr(a[1]), _hptr(a[2])))

    def xchg_views(self, side, nrows, nbytes):
        """(row ids int32[nrows], lengths int32[nrows], key bytes uint8[nbytes]) of this rank's
        receive buffers, as tensors over the library's memory"""
        t = self._torch
        pc, pr, pl = C.c_void_p(), C.c_void_p(), C.c_void_p()
        rc, bc = C.c_size_t(), C.c_size_t()
        self._ck(self.lib.hmj_xchg_buffers(self.h, side, C.byref(pc), C.byref(pr), C.byref(pl), C.byref(rc), C.byref(bc)))
        assert nrows <= rc.value and nbytes <= bc.value
        def view(ptr, n, typestr, dtype):
            if n == 0:
                return t.empty(0, dtype=dtype, device=self.device)
            return t.as_tensor(_DeviceSpan(ptr.value, n, typestr), device=self.device)
        return (view(pr, nrows, "<i4", t.int32), view(pl, nrows, "<i4", t.int32), view(pc, nbytes, "|u1", t.uint8))

    def xchg_teardown(self):
        self._ck(self.lib.hmj_xchg_teardown(self.h))

    # ---- device generators of the benchmark relations (datagen.cu)
    def gen_permute(self, n, domain, seed, src=None, base=0):
        """int32 tensor: P(src[i]) or P(base + i), P a seeded bijection of [0, domain)"""
        out = self._empty(n, self._torch.int32)
        self._ck(self.lib.hmj_gen_permute(self.h, _dptr(src), base, n, domain, seed & 0xFFFFFFFF, _dptr(out)))
        return out

    def gen_zipf(self, cdf, n, seed, row_base=0):
        """int32 tensor of n ranks drawn from the float64 device prefix table `cdf`"""
        assert cdf.is_cuda and cdf.dtype == self._torch.float64 and cdf.is_contiguous()
        out = self._empty(n, self._torch.int32)
        self._ck(self.lib.hmj_gen_zipf(self.h, _dptr(cdf), cdf.numel(), seed, row_base, n, _dptr(out)))
        return out

    def gen_strgen(self, word_offsets, word_chars, key_ids):
        """(offsets int64[n+1], chars uint8[]) of the strgen keys `key_ids` (int32 tensor)"""
        t = self._torch
        n, q = key_ids.numel(), word_offsets.numel() - 1
        lens = self._empty(n, t.int64)
        self._ck(self.lib.hmj_gen_strgen_lengths(self.h, _dptr(word_offsets), q, _dptr(key_ids), n, _dptr(lens)))
        off = t.zeros(n + 1, dtype=t.int64, device=self.device)
        if n:
            t.cumsum(lens, 0, out=off[1:])
        del lens
        total = int(off[-1].item()) if n else 0
        chars = self._empty(total + 16, t.uint8)[:total]   # padded: readers may look one word ahead
        self._ck(self.lib.hmj_gen_strgen_write(self.h, _dptr(word_offsets), _dptr(word_chars), q,
                                               _dptr(key_ids), n, _dptr(off), _dptr(chars)))
        return off, chars

    # ---- misc
    def synchronize(self):
        self._ck(self.lib.hmj_synchronize(self.h))

    def launch_count(self):
        c = C.c_uint64()
        self._ck(self.lib.hmj_launch_count(self.h, C.byref(c)))
        return c.value

    def set_profiling(self, on):
        self._ck(self.lib.hmj_set_profiling(self.h, 1 if on else 0))

    def timings(self):
        t = Timings()
        self._ck(self.lib.hmj_get_timings(self.h, C.byref(t)))
        out = {k: getattr(t, k) for k, _ in Timings._fields_}
        for k in ("kernel_launches", "kernel_ms", "kernel_bytes"):
            out[k] = list(out[k])
        return out


def optimal_partition(n):
    """radix_hash::optimal_partition (radix_hash.h:38-57); host scalar, works without a GPU."""
    return _lib.load().hmj_optimal_partition(int(n))
