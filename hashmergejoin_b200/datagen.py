"""Seeded synthetic relations for tests and benchmarks (harness code, not part of the hot path).

* ``strgen``  -- relations shaped like the reference's ``create_strvec`` (/root/reference
  strgen.cc:28-61): the first q = ceil(sqrt(n)) rows are base words, the rest are
  ``word_i + "-" + word_j`` row-major until n rows, then shuffled.  The reference reads
  /usr/share/dict/words, which does not exist in this image, so base words come from a seeded
  synthetic list of unique lowercase words (lengths 3..14, mean ~8.5, row mean ~18 bytes).
* ``unique_u64`` -- n distinct uniform-looking 64-bit keys (splitmix64 finaliser of 0..n-1,
  which is a bijection on 64 bits), and ``permuted`` -- a seeded permutation of a relation.
Columnar layout: offsets u64[n+1] + chars u8[] for strings, u64[n] for integers.
"""
import math

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """Bijective 64-bit mix (numpy uint64, wrapping arithmetic)."""
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return x


def unique_u64(n, seed=0):
    """n distinct 64-bit keys."""
    base = np.arange(n, dtype=np.uint64) + (np.uint64(seed) << np.uint64(40))
    return splitmix64(base)


def permuted(arr, seed):
    rng = np.random.default_rng(seed)
    return arr[rng.permutation(arr.shape[0])]


def word_list(q, seed=1234):
    """q unique lowercase words of length 3..14 as (offsets u64[q+1], chars u8[])."""
    rng = np.random.default_rng(seed)
    words = set()
    out = []
    while len(out) < q:
        need = q - len(out)
        lens = rng.integers(3, 15, size=need * 2)
        for ln in lens:
            w = bytes(rng.integers(97, 123, size=int(ln), dtype=np.uint8))
            if w not in words:
                words.add(w)
                out.append(w)
                if len(out) == q:
                    break
    lens = np.array([len(w) for w in out], dtype=np.uint64)
    offsets = np.zeros(q + 1, dtype=np.uint64)
    np.cumsum(lens, out=offsets[1:])
    chars = np.frombuffer(b"".join(out), dtype=np.uint8).copy()
    return offsets, chars


def strgen_unshuffled(n, seed=1234):
    """create_strvec(n) before the shuffle: (offsets, chars, values)."""
    q = int(math.ceil(math.sqrt(n)))
    woff, wchars = word_list(q, seed)
    wlen = (woff[1:] - woff[:-1]).astype(np.int64)
    nbase = min(q, n)
    npairs = n - nbase
    # row r >= q is pair (i, j) = divmod(r - q, q)
    pi = np.arange(npairs, dtype=np.int64) // q
    pj = np.arange(npairs, dtype=np.int64) % q
    lens = np.concatenate([wlen[:nbase], wlen[pi] + 1 + wlen[pj]]).astype(np.uint64)
    offsets = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=offsets[1:])
    chars = np.empty(int(offsets[-1]), dtype=np.uint8)
    # base words
    chars[: int(offsets[nbase])] = wchars[: int(woff[nbase])]
    # pairs, filled by (vectorised) per-length-class gathers
    if npairs:
        start = offsets[nbase:-1].astype(np.int64)
        li, lj = wlen[pi], wlen[pj]
        wo = woff[:-1].astype(np.int64)
        for k in range(int(wlen.max())):
            sel = li > k
            chars[start[sel] + k] = wchars[wo[pi[sel]] + k]
            sel = lj > k
            chars[start[sel] + li[sel] + 1 + k] = wchars[wo[pj[sel]] + k]
        chars[start + li] = ord("-")
    values = np.concatenate([np.arange(nbase, dtype=np.uint64), (pi + pj).astype(np.uint64)])
    return offsets, chars, values


def take_rows(offsets, chars, order):
    """Reorders a string column: row k of the result is row order[k] of the input."""
    lens = (offsets[1:] - offsets[:-1])[order]
    new_off = np.zeros(order.shape[0] + 1, dtype=np.uint64)
    np.cumsum(lens, out=new_off[1:])
    src_start = offsets[:-1][order].astype(np.int64)
    # index of every output byte: repeat starts and add intra-row position
    total = int(new_off[-1])
    row_of_byte = np.repeat(np.arange(order.shape[0], dtype=np.int64), lens.astype(np.int64))
    pos = np.arange(total, dtype=np.int64) - new_off[:-1].astype(np.int64)[row_of_byte]
    new_chars = chars[src_start[row_of_byte] + pos]
    return new_off, new_chars


def strgen(n, seed=1234, shuffle_seed=1):
    """create_strvec(n): unique keys, shuffled with a seeded permutation.
    Returns (offsets, chars, values, order) where order[k] = unshuffled row of output row k."""
    offsets, chars, values = strgen_unshuffled(n, seed)
    order = np.random.default_rng(shuffle_seed).permutation(n)
    new_off, new_chars = take_rows(offsets, chars, order)
    return new_off, new_chars, values[order], order


def strings_from_list(items):
    """Python list of bytes/str -> (offsets, chars)."""
    bs = [s.encode() if isinstance(s, str) else bytes(s) for s in items]
    offsets = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        np.cumsum(np.array([len(b) for b in bs], dtype=np.uint64), out=offsets[1:])
    chars = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, np.uint8)
    if chars.size == 0:
        chars = np.zeros(1, np.uint8)[:0].copy()
    return offsets, chars


def zipf_indices(n_draws, n_keys, s=1.0, seed=7):
    """n_draws ranks in [0,n_keys) with P(k) ~ 1/(k+1)^s via a seeded inverse CDF."""
    rng = np.random.default_rng(seed)
    w = 1.0 / np.power(np.arange(1, n_keys + 1, dtype=np.float64), s)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    u = rng.random(n_draws)
    return np.minimum(np.searchsorted(cdf, u), n_keys - 1).astype(np.int64)


# ---- numpy restatements of the device generators in csrc/datagen.cu ----------------------
def _fmix32(h):
    h = h.astype(np.uint32)
    h ^= h >> np.uint32(16)
    h = (h.astype(np.uint64) * np.uint64(0x85EBCA6B)).astype(np.uint32)
    h ^= h >> np.uint32(13)
    h = (h.astype(np.uint64) * np.uint64(0xC2B2AE35)).astype(np.uint32)
    h ^= h >> np.uint32(16)
    return h


def feistel_permute(x, domain, seed):
    """hmj_gen_permute: a seeded bijection of [0, domain) applied to the uint32 array x
    (4-round balanced Feistel network on the next even number of bits, cycle-walked)."""
    half_bits = 1
    while half_bits < 16 and (1 << (2 * half_bits)) < domain:
        half_bits += 1
    hm = np.uint32((1 << half_bits) - 1)
    x = np.asarray(x, dtype=np.uint32).copy()
    todo = np.ones(x.shape, bool)
    while todo.any():
        v = x[todo]
        l, r = v >> np.uint32(half_bits), v & hm
        for rnd in range(4):
            key = np.uint32((seed + rnd * 0x9E3779B9) & 0xFFFFFFFF)
            f = _fmix32(r ^ key) & hm
            l, r = r, l ^ f
        v = (l << np.uint32(half_bits)) | r
        x[todo] = v
        todo[todo] = v >= domain
    return x


def zipf_cdf(n_keys, s=1.0):
    """float64 prefix table of P(k) ~ 1/(k+1)^s, last entry exactly 1.0"""
    w = 1.0 / np.power(np.arange(1, n_keys + 1, dtype=np.float64), s)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    cdf[-1] = 1.0
    return cdf


def zipf_ranks(cdf, n, seed, row_base=0):
    """hmj_gen_zipf: rank of row i = first k with cdf[k] > u(seed, row_base + i)"""
    with np.errstate(over="ignore"):
        ctr = np.uint64(seed) * np.uint64(0xD1342543DE82EF95) + np.uint64(row_base) + np.arange(n, dtype=np.uint64)
    u = (splitmix64(ctr) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    return np.minimum(np.searchsorted(cdf, u, side="right"), cdf.size - 1).astype(np.uint32)


def strgen_keys(woff, wchars, key_ids):
    """hmj_gen_strgen_*: (offsets, chars) of the strgen keys `key_ids` over the word table:
    id < q is word id, otherwise word_i + "-" + word_j with (i, j) = divmod(id - q, q)."""
    q = woff.size - 1
    out = []
    for k in np.asarray(key_ids, dtype=np.int64):
        w = lambda i: bytes(wchars[int(woff[i]):int(woff[i + 1])])
        out.append(w(k) if k < q else w((k - q) // q) + b"-" + w((k - q) % q))
    return strings_from_list(out)


class DeviceStrgen:
    """strgen-shaped relations generated in HBM (csrc/datagen.cu) over `n_keys` distinct keys:
    the word table is the host's seeded synthetic list (word_list), everything per row runs on
    the GPU.  Relations are described by a key id per row."""

    def __init__(self, ctx, n_keys, seed=1234):
        import torch
        self.ctx, self.n_keys = ctx, n_keys
        self.q = int(math.ceil(math.sqrt(n_keys)))
        woff, wchars = word_list(self.q, seed)
        self.h_woff, self.h_wchars = woff, wchars
        self.woff = torch.from_numpy(woff.view(np.int64)).to(ctx.device)
        self.wchars = torch.from_numpy(wchars).to(ctx.device)

    def shuffled_ids(self, n_rows, seed, copies=1, row_base=0):
        """key ids of rows [row_base, row_base + n_rows) of a relation holding every key `copies`
        times in a seeded random order (copies = 1: create_strvec's shuffle)"""
        ids = self.ctx.gen_permute(n_rows, self.n_keys * copies, seed, base=row_base)
        if copies > 1:
            ids = (ids // copies).to(ids.dtype)   # values < 2^31: plain integer division
        return ids

    def zipf_ids(self, cdf, n_rows, seed, key_seed, row_base=0):
        """Zipf-distributed key ids: rank k -> key P(k), so that the hot keys are arbitrary ones"""
        ranks = self.ctx.gen_zipf(cdf, n_rows, seed, row_base)
        return self.ctx.gen_permute(n_rows, self.n_keys, key_seed, src=ranks)

    def keys(self, ids):
        return self.ctx.gen_strgen(self.woff, self.wchars, ids)


# ---- device-side generators (torch; harness only) for the full-size configurations -------
def _lsr(x, s):
    """logical shift right on int64 tensors (torch's >> is arithmetic)"""
    return (x >> s) & ((1 << (64 - s)) - 1)


def splitmix64_torch(x):
    """Same bijection as splitmix64() on int64 tensors holding u64 bit patterns."""
    import torch
    c1 = torch.tensor(-7046029254386353131, dtype=torch.int64, device=x.device)   # 0x9E3779B97F4A7C15
    c2 = torch.tensor(-4658895280553007687, dtype=torch.int64, device=x.device)   # 0xBF58476D1CE4E5B9
    c3 = torch.tensor(-7723592293110705685, dtype=torch.int64, device=x.device)   # 0x94D049BB133111EB
    x = x + c1
    x = (x ^ _lsr(x, 30)) * c2
    x = (x ^ _lsr(x, 27)) * c3
    return x ^ _lsr(x, 31)


def unique_u64_torch(n, seed=0, device="cuda", perm_mult=None, chunk=1 << 26):
    """Device version of unique_u64: key[i] = splitmix64(idx(i) + (seed<<40)) with
    idx(i) = i, or a permutation of 0..n-1 when perm_mult is given: idx(i) = (i*perm_mult + 12345)
    mod n for n a power of two and odd perm_mult (a bijection), so two calls with different
    perm_mult give the same key set in different row orders."""
    import torch
    out = torch.empty(n, dtype=torch.int64, device=device)
    if perm_mult is not None:
        assert n & (n - 1) == 0 and perm_mult % 2 == 1, "permutation needs n = 2^k and an odd multiplier"
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        idx = torch.arange(lo, hi, dtype=torch.int64, device=device)
        if perm_mult is not None:
            idx = (idx * perm_mult + 12345) & (n - 1)
        out[lo:hi] = splitmix64_torch(idx + (seed << 40))
    return out
