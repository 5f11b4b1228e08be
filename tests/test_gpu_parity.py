"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same
seeded inputs, against the committed golden fixtures, and -- at full benchmark sizes -- through
size-independent properties.  Bit-exact everywhere (integer/byte/index work)."""
import numpy as np
import pytest

import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import datagen
from tests import golden_util as G
from tests.oracle_binding import Keys

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    """numpy u64/u32/u8 -> device tensor with the same bits"""
    a = np.ascontiguousarray(a)
    view = {np.dtype(np.uint64): np.int64, np.dtype(np.uint32): np.int32}.get(a.dtype, None)
    t = torch.from_numpy(a.view(view) if view else a)
    return t.cuda()


def host(t, dt):
    return t.cpu().numpy().view(dt)


# ------------------------------------------------------------------ K1 hashing
def test_hash_golden_vectors(ctx):
    g = G.load("std_hash_strings.json")
    off, chars = datagen.strings_from_list([bytes.fromhex(s) for s in g["strings_hex"]])
    chars = np.concatenate([chars, np.zeros(8, np.uint8)])
    out = host(ctx.hash_strings(dev(off), dev(chars)), np.uint64)
    assert np.array_equal(out, G.u64(g["hashes"]))


def test_hash_random_all_lengths_and_alignments(ctx, oracle):
    rng = np.random.default_rng(1)
    items = [bytes(rng.integers(0, 256, size=ln, dtype=np.uint8)) for ln in range(0, 100) for _ in range(40)]
    rng.shuffle(items)
    off, chars = datagen.strings_from_list(items)
    out = host(ctx.hash_strings(dev(off), dev(chars)), np.uint64)
    assert np.array_equal(out, oracle.hash_strings(off, chars))
    # same rows behind a 3-byte prefix: every row starts at another alignment
    off3 = off + np.uint64(3)
    chars3 = np.concatenate([np.array([1, 2, 3], np.uint8), chars])
    out3 = host(ctx.hash_strings(dev(off3), dev(chars3)), np.uint64)
    assert np.array_equal(out3, out)


def test_hash_strgen_and_long_rows(ctx, oracle):
    off, chars, _, _ = datagen.strgen(200000, seed=3)
    out = host(ctx.hash_strings(dev(off), dev(chars)), np.uint64)
    assert np.array_equal(out, oracle.hash_strings(off, chars))
    # rows long enough that a 1024-row tile cannot be staged in shared memory
    rng = np.random.default_rng(2)
    items = [bytes(rng.integers(0, 256, size=int(ln), dtype=np.uint8)) for ln in rng.integers(0, 400, size=3000)]
    off, chars = datagen.strings_from_list(items)
    out = host(ctx.hash_strings(dev(off), dev(chars)), np.uint64)
    assert np.array_equal(out, oracle.hash_strings(off, chars))


def test_hash_pipelined_kernel_regular_and_irregular_tiles(ctx, oracle):
    # >= 4 tiles of 1024 rows: the bulk-copy pipelined kernel.  Short rows (tiles staged in shared
    # memory), long rows (tiles whose bytes do not fit: hashed from global memory), a mix of both in
    # one column, empty rows at the end, and a row count that leaves a partial last tile.
    rng = np.random.default_rng(5)
    def rows(lens):
        return [bytes(rng.integers(0, 256, size=int(ln), dtype=np.uint8)) for ln in lens]
    cases = {
        "short": rows(rng.integers(0, 30, size=9000)),
        "long": rows(rng.integers(0, 400, size=6001)),
        "mixed": rows(np.concatenate([rng.integers(0, 20, size=3000), rng.integers(100, 300, size=2500),
                                      rng.integers(0, 20, size=3000), np.zeros(1500, np.int64)])),
    }
    for name, items in cases.items():
        off, chars = datagen.strings_from_list(items)
        chars = np.concatenate([chars, np.zeros(64, np.uint8)])
        out = host(ctx.hash_strings(dev(off), dev(chars)), np.uint64)
        assert np.array_equal(out, oracle.hash_strings(off, chars)), name


def test_hash_u64_identity(ctx):
    k = np.random.default_rng(0).integers(0, 2 ** 64, size=100001, dtype=np.uint64)
    assert np.array_equal(host(ctx.hash_u64(dev(k)), np.uint64), k)


# ------------------------------------------------------------------ K2/K3 sort
def check_sort(ctx, oracle, keys, rowid=None):
    keys = np.ascontiguousarray(keys, np.uint64)
    h, r = ctx.sort_pairs(dev(keys), None if rowid is None else dev(rowid))
    h, r = host(h, np.uint64), host(r, np.uint32)
    rid = np.arange(keys.size, dtype=np.uint32) if rowid is None else rowid
    # (1) hash column == reference order (== std::sort of the hashes)
    oh, orow = oracle.radix_sort(keys, Keys(u64=keys), rowid=rid)
    assert np.array_equal(h, oh)
    # (2) same (hash,row) multiset; our sort is stable so it IS the canonical order
    ch, cr = oracle.canonical(oh, orow)
    if rowid is None:
        assert np.array_equal(r, cr)
    else:
        assert np.array_equal(keys[np.argsort(keys, kind="stable")], h)
        assert np.array_equal(rid[np.argsort(keys, kind="stable")], r)


@pytest.mark.parametrize("case", [np.arange(4, -1, -1), np.arange(5, 0, -1), np.arange(12345, 0, -1),
                                  np.arange(1 << 18, 0, -1)])
def test_sort_reference_ladder(ctx, oracle, case):
    # radix_hash_test.cc:40-120 (+ in-place variants :122-193): reversed ranges
    check_sort(ctx, oracle, case.astype(np.uint64))


@pytest.mark.parametrize("n", [1, 2, 31, 4095, 4096, 4097, 1 << 16, 100003, (1 << 20) + 5])
def test_sort_random_u64(ctx, oracle, n):
    # radix_hash_test.cc:195-212,244-262 / radix_sort_test.cc:27-68: random u64 vs std::sort
    check_sort(ctx, oracle, np.random.default_rng(n).integers(0, 2 ** 64, size=n, dtype=np.uint64))


def test_sort_top_bit_and_duplicates(ctx, oracle):
    top = np.empty(1024, np.uint64)  # radix_hash_test.cc:227-242
    top[0::2] = (np.uint64(1) << np.uint64(63)) + np.arange(512, dtype=np.uint64)
    top[1::2] = np.arange(512, dtype=np.uint64)
    check_sort(ctx, oracle, top)
    rng = np.random.default_rng(9)
    check_sort(ctx, oracle, rng.integers(0, 7, size=50000, dtype=np.uint64))          # heavy duplicates
    check_sort(ctx, oracle, np.full(9000, 0xFFFFFFFFFFFFFFFF, np.uint64))             # all max
    check_sort(ctx, oracle, rng.integers(0, 2 ** 64, size=30000, dtype=np.uint64) | np.uint64(0xFF))
    vals = rng.permutation(70000).astype(np.uint32)                                   # explicit payloads
    check_sort(ctx, oracle, rng.integers(0, 1000, size=70000, dtype=np.uint64), rowid=vals)


def test_sort_skips_trivial_passes_large(ctx, oracle):
    n = (1 << 22) + 17  # above the skip threshold: keys < 2^20 -> only 3 passes run
    keys = np.random.default_rng(5).integers(0, 1 << 20, size=n, dtype=np.uint64)
    check_sort(ctx, oracle, keys)
    # in place (input aliases output) with an odd number of executed passes
    t = dev(keys)
    r = torch.empty(n, dtype=torch.int32, device="cuda")
    ctx.sort_pairs(t, None, out_hash=t, out_rowid=r)
    assert np.array_equal(host(t, np.uint64), np.sort(keys))
    assert np.array_equal(keys[host(r, np.uint32)], np.sort(keys))


def fast_check_sort(ctx, keys, rowid=None, inplace=False):
    """large cases: numpy stable argsort is the check (== std::sort order of the reference tests)"""
    keys = np.ascontiguousarray(keys, np.uint64)
    t = dev(keys)
    if inplace:
        r = torch.empty(keys.size, dtype=torch.int32, device="cuda")
        ctx.sort_pairs(t, None if rowid is None else dev(rowid), out_hash=t, out_rowid=r)
        h = t
    else:
        h, r = ctx.sort_pairs(t, None if rowid is None else dev(rowid))
    order = np.argsort(keys, kind="stable")
    rid = np.arange(keys.size, dtype=np.uint32) if rowid is None else rowid
    assert np.array_equal(host(h, np.uint64), keys[order])
    assert np.array_equal(host(r, np.uint32), rid[order])
    return ctx.timings()


def test_sort_hybrid_prefix_and_bucket_finish(ctx, monkeypatch):
    monkeypatch.setenv("HMJ_SORT_UNIFORM", "0")  # the histogram-planned path is the subject here
    # uniform keys, n >= 2^21: 3 scatter passes on the top 24 bits + bucket finish (shift 40)
    rng = np.random.default_rng(21)
    n = (1 << 21) + 12345
    keys = rng.integers(0, 2 ** 64, size=n, dtype=np.uint64)
    f0 = ctx.timings()["sort_finishes"]
    tm = fast_check_sort(ctx, keys)
    assert tm["sort_finishes"] == f0 + 1, "the hybrid path did not run"
    fast_check_sort(ctx, keys, inplace=True)
    fast_check_sort(ctx, keys, rowid=rng.permutation(n).astype(np.uint32))
    # every key twice: equal keys inside a bucket keep their input order
    fast_check_sort(ctx, np.concatenate([keys[: n // 2], keys[: n // 2]]))
    # 2^17 rows: two scatter passes, finish on shift 48
    fast_check_sort(ctx, rng.integers(0, 2 ** 64, size=1 << 17, dtype=np.uint64))


def test_sort_hybrid_tie_slow_path(ctx, monkeypatch):
    monkeypatch.setenv("HMJ_SORT_UNIFORM", "0")  # the histogram-planned path is the subject here
    # bucket-mates that agree on the 31 bits below the 24-bit prefix and differ only in the low
    # 9 bits: the 32-bit fast ranking cannot order them, the tile must take the full-key path
    rng = np.random.default_rng(22)
    n = 1 << 21
    top = rng.integers(0, 1 << 20, size=n, dtype=np.uint64) << np.uint64(44)
    low = rng.integers(0, 512, size=n, dtype=np.uint64)
    tm0 = ctx.timings()
    tm = fast_check_sort(ctx, top | low)
    assert tm["sort_finishes"] == tm0["sort_finishes"] + 1
    assert tm["sort_fallbacks"] == tm0["sort_fallbacks"]
    # mixed: only some buckets tie
    keys = rng.integers(0, 2 ** 64, size=n, dtype=np.uint64)
    keys[::7] = (keys[::7] & ~np.uint64((1 << 40) - 1)) | low[::7]
    fast_check_sort(ctx, keys)


def test_sort_hybrid_oversized_bucket_falls_back(ctx, monkeypatch):
    monkeypatch.setenv("HMJ_SORT_UNIFORM", "0")  # the histogram-planned path is the subject here
    # digit histograms look uniform (1200 extra rows in a bin of 32768 is inside the planner's
    # tolerance), but 1200 rows share one 24-bit prefix: the bucket finish flags it and the sort
    # re-runs with scatter passes over every digit -- same result
    rng = np.random.default_rng(23)
    n = 1 << 23
    keys = rng.integers(0, 2 ** 64, size=n, dtype=np.uint64)
    hot = rng.choice(n, size=1200, replace=False)
    keys[hot] = (np.uint64(0xABCDEF) << np.uint64(40)) | rng.integers(0, 1 << 40, size=1200, dtype=np.uint64)
    tm0 = ctx.timings()
    tm = fast_check_sort(ctx, keys)
    assert tm["sort_fallbacks"] == tm0["sort_fallbacks"] + 1
    fast_check_sort(ctx, keys, inplace=True)
    # 1200 copies of ONE key (a moderate heavy hitter the histogram test cannot see)
    keys[hot] = np.uint64(0x123456789ABCDEF0)
    fast_check_sort(ctx, keys)


# ---- uniform-hash path (bucket_scatter.cu + padded segment sort): iota row ids, no histogram
@pytest.mark.parametrize("n", [4609, 5000, 100003, (1 << 20) + 5, (1 << 21) + 12345, 1 << 23])
def test_sort_uniform_path_random(ctx, n):
    rng = np.random.default_rng(n)
    keys = rng.integers(0, 2 ** 64, size=n, dtype=np.uint64)
    u0 = ctx.timings()["sort_uniform"]
    tm = fast_check_sort(ctx, keys)
    assert tm["sort_uniform"] == u0 + 1, "the uniform path did not run"


def test_sort_uniform_path_duplicates_are_ordered_by_row_id(ctx):
    rng = np.random.default_rng(31)
    n = 1 << 21
    base = rng.integers(0, 2 ** 64, size=n // 4, dtype=np.uint64)
    keys = np.concatenate([base, base, base, base])[rng.permutation(n)]  # every key four times
    u0 = ctx.timings()["sort_uniform"]
    tm = fast_check_sort(ctx, keys)
    assert tm["sort_uniform"] == u0 + 1
    # runs of 40 equal keys
    base = rng.integers(0, 2 ** 64, size=n // 40, dtype=np.uint64)
    keys = np.repeat(base, 40)[rng.permutation(base.size * 40)]
    fast_check_sort(ctx, keys)
    # keys that differ only in their low bits inside a bucket
    top = rng.integers(0, 1 << 20, size=n, dtype=np.uint64) << np.uint64(44)
    fast_check_sort(ctx, top | rng.integers(0, 512, size=n, dtype=np.uint64))


def test_sort_uniform_path_declines_or_recovers(ctx):
    rng = np.random.default_rng(32)
    n = 1 << 22
    # skewed: everything in one top-level bucket -> overflow flag -> histogram-planned sort
    u0 = ctx.timings()["sort_uniform"]
    tm = fast_check_sort(ctx, rng.integers(0, 1 << 40, size=n, dtype=np.uint64))
    assert tm["sort_uniform"] == u0
    # one final bucket too full (1200 extra rows share a 24-bit prefix)
    keys = rng.integers(0, 2 ** 64, size=n, dtype=np.uint64)
    hot = rng.choice(n, size=1200, replace=False)
    keys[hot] = (np.uint64(0xABCDEF) << np.uint64(40)) | rng.integers(0, 1 << 40, size=1200, dtype=np.uint64)
    fast_check_sort(ctx, keys)
    # 1200 copies of one key: a crowded bin inside the bucket
    keys[hot] = np.uint64(0x123456789ABCDEF0)
    fast_check_sort(ctx, keys)
    # explicit payloads never take the uniform path (it orders ties by row id)
    u0 = ctx.timings()["sort_uniform"]
    tm = fast_check_sort(ctx, rng.integers(0, 2 ** 64, size=n, dtype=np.uint64),
                         rowid=rng.permutation(n).astype(np.uint32))
    assert tm["sort_uniform"] == u0


def test_sort_hybrid_small_whole_array_bucket(ctx):
    rng = np.random.default_rng(24)
    for n in (3, 200, 256, 257, 511, 512, 513):
        fast_check_sort(ctx, rng.integers(0, 2 ** 64, size=n, dtype=np.uint64))
        fast_check_sort(ctx, rng.integers(0, 5, size=n, dtype=np.uint64), inplace=True)


def test_sort_empty(ctx):
    h, r = ctx.sort_pairs(torch.empty(0, dtype=torch.int64, device="cuda"))
    assert h.numel() == 0 and r.numel() == 0


def test_hash_sort_strings_golden_and_oracle(ctx, oracle):
    g = G.load("strgen_join_600.json")
    off, chars = np.array(g["r_offsets"], np.uint64), G.u8(g["r_chars_hex"])
    h, r = ctx.hash_sort_strings(dev(off), dev(np.concatenate([chars, np.zeros(8, np.uint8)])))
    assert np.array_equal(host(h, np.uint64), G.u64(g["r_sorted_hash"]))   # reference's sorted hashes
    assert np.array_equal(host(r, np.uint32), np.array(g["r_sorted_rowid"], np.uint32))  # unique hashes
    off, chars, _, _ = datagen.strgen(300000, seed=4)
    h, r = ctx.hash_sort_strings(dev(off), dev(chars))
    k = Keys(offsets=off, chars=chars)
    oh, orow = oracle.canonical(*oracle.radix_sort(oracle.hashes(k), k))
    assert np.array_equal(host(h, np.uint64), oh) and np.array_equal(host(r, np.uint32), orow)


# ------------------------------------------------------------------ K4-K6 merge join
def gpu_join_u64(ctx, rk, sk):
    rh, rr = ctx.hash_sort_u64(dev(rk))
    sh, sr = ctx.hash_sort_u64(dev(sk))
    R, S = hmj.SortedRelation(rh, rr), hmj.SortedRelation(sh, sr)
    out_r, out_s = ctx.merge_join(R, S)
    m, sm = ctx.merge_join_aggregate(R, S)
    return host(out_r, np.uint32), host(out_s, np.uint32), m, sm


def check_join_u64(ctx, oracle, rk, sk):
    rk, sk = np.ascontiguousarray(rk, np.uint64), np.ascontiguousarray(sk, np.uint64)
    out_r, out_s, m, sm = gpu_join_u64(ctx, rk, sk)
    t_r, t_s, t_sum = oracle.full_join(Keys(u64=rk), Keys(u64=sk))
    c_r, c_s = oracle.sort_pairs(out_r, out_s)
    assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)
    assert m == t_r.size and sm == t_sum
    return out_r, out_s


def test_join_unique_u64(ctx, oracle):
    keys = datagen.unique_u64(100000, seed=1)
    out_r, _ = check_join_u64(ctx, oracle, datagen.permuted(keys, 1), datagen.permuted(keys, 2))
    assert out_r.size == 100000


def check_operator_u64(ctx, oracle, rk, sk):
    """the whole operator (hmj_join_u64: HashMergeJoin ctor + full iteration) against the oracle"""
    rk, sk = np.ascontiguousarray(rk, np.uint64), np.ascontiguousarray(sk, np.uint64)
    m = ctx.join_u64(dev(rk), dev(sk))
    out_r, out_s = ctx.join_result_to_host(m)
    t_r, t_s, _ = oracle.full_join(Keys(u64=rk), Keys(u64=sk))
    c_r, c_s = oracle.sort_pairs(out_r, out_s)
    assert m == t_r.size and np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)
    return m


def test_operator_bucket_join_path(ctx, oracle):
    # uniform hashes: the whole-operator join takes the bucket path (bucket_join.cu)
    rng = np.random.default_rng(41)
    n = 900000
    keys = datagen.unique_u64(n, seed=5)
    rk, sk = datagen.permuted(keys, 1).copy(), datagen.permuted(keys, 2).copy()
    assert check_operator_u64(ctx, oracle, rk, sk) == n
    assert ctx.timings()["bucket_join"] == 1, "the bucket join did not run"
    assert ctx.timings()["scatter_levels"] == 2, "one scatter level per relation at this size"
    # 300 x 300 copies of one key: 90000 pairs from one bucket (more than its staging area)
    rk[rng.choice(n, 300, replace=False)] = np.uint64(0x9E3779B97F4A7C15)
    sk[rng.choice(n, 300, replace=False)] = np.uint64(0x9E3779B97F4A7C15)
    assert check_operator_u64(ctx, oracle, rk, sk) >= 90000
    assert ctx.timings()["bucket_join"] == 1
    # different sizes on the two sides share the bucket numbering of the larger one
    check_operator_u64(ctx, oracle, rk[:50000], sk)
    check_operator_u64(ctx, oracle, rk, sk[:7000])
    # every key twice on one side, four times on the other, plus non-matching rows
    base = datagen.unique_u64(120000, seed=6)
    rk2 = np.concatenate([base[:100000], base[:100000]])[rng.permutation(200000)]
    sk2 = np.concatenate([base[20000:]] * 4)[rng.permutation(400000)]
    check_operator_u64(ctx, oracle, rk2, sk2)


@pytest.mark.parametrize("nr,ns,rng_hi", [(0, 10, 5), (10, 0, 5), (1, 3, 1), (5000, 7000, 600),
                                          (30000, 100, 50), (20000, 20000, 2 ** 63), (4609, 4609, 2 ** 64)])
def test_operator_shapes_fall_back_or_not(ctx, oracle, nr, ns, rng_hi):
    rng = np.random.default_rng(nr * 11 + ns)
    check_operator_u64(ctx, oracle, rng.integers(0, rng_hi, size=nr, dtype=np.uint64),
                       rng.integers(0, rng_hi, size=ns, dtype=np.uint64))


def test_join_pairs_of_one_hash_range(ctx, oracle):
    # the local step of the multi-GPU join: unsorted (hash, global row id) pairs whose hashes share
    # their top bits (rank 2 of 4)
    rng = np.random.default_rng(43)
    n = 700000
    keys = (datagen.unique_u64(n, seed=9) >> np.uint64(2)) | (np.uint64(2) << np.uint64(62))
    rk, sk = datagen.permuted(keys, 1).copy(), datagen.permuted(keys, 2)[: n // 2].copy()
    rk[:1000] = rk[1000:2000]  # some duplicates on the build side
    rid = (rng.permutation(n) + 5 * n).astype(np.uint32)
    sid = (rng.permutation(sk.size) + 9 * n).astype(np.uint32)
    for bits in (2, 0):
        m = ctx.join_pairs(dev(rk), dev(rid), dev(sk), dev(sid), prefix_bits=bits)
        # prefix 0 declares no shared bits: the range fills a quarter of the buckets, they overflow
        # and the stable sort + merge join takes over -- same result
        assert ctx.timings()["bucket_join"] == (1 if bits == 2 else 0)
        out_r, out_s = ctx.join_result_to_host(m)
        t_r, t_s, _ = oracle.full_join(Keys(u64=rk), Keys(u64=sk))
        e_r, e_s = oracle.sort_pairs(rid[t_r], sid[t_s])
        c_r, c_s = oracle.sort_pairs(out_r, out_s)
        assert np.array_equal(c_r, e_r) and np.array_equal(c_s, e_s)
    # skewed range (everything in one bucket): stable sort + merge join
    rk2 = (rng.integers(0, 1 << 20, size=50000, dtype=np.uint64)) | (np.uint64(2) << np.uint64(62))
    sk2 = (rng.integers(0, 1 << 20, size=60000, dtype=np.uint64)) | (np.uint64(2) << np.uint64(62))
    m = ctx.join_pairs(dev(rk2), dev(rid[:50000].copy()), dev(sk2), dev(sid[:60000].copy()), prefix_bits=2)
    assert ctx.timings()["bucket_join"] == 0
    out_r, out_s = ctx.join_result_to_host(m)
    t_r, t_s, _ = oracle.full_join(Keys(u64=rk2), Keys(u64=sk2))
    e_r, e_s = oracle.sort_pairs(rid[:50000][t_r], sid[:60000][t_s])
    c_r, c_s = oracle.sort_pairs(out_r, out_s)
    assert np.array_equal(c_r, e_r) and np.array_equal(c_s, e_s)


def test_operator_strings_bucket_join(ctx, oracle):
    off, chars, _, _ = datagen.strgen(50000, seed=7, shuffle_seed=1)
    rng = np.random.default_rng(4)
    roff, rchars = datagen.take_rows(off, chars, rng.integers(0, 50000, size=60000))
    soff, schars = datagen.take_rows(off, chars, rng.integers(0, 50000, size=40000))
    out_r, out_s = gpu_join_strings(ctx, roff, rchars, soff, schars)
    assert ctx.timings()["bucket_join"] == 1, "the bucket join did not run"
    t_r, t_s, _ = oracle.full_join(Keys(offsets=roff, chars=rchars), Keys(offsets=soff, chars=schars))
    c_r, c_s = oracle.sort_pairs(out_r, out_s)
    assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)


@pytest.mark.parametrize("skewed_side", ["probe", "build"])
def test_operator_zipf_relation_trades_places(ctx, oracle, skewed_side):
    # C5's shape: one relation holds every key twice, the other draws keys Zipf(1.0); the head
    # keys' runs span many tiles, so the join iterates over the skewed relation whichever side
    # it is on -- the pairs must not change
    dkeys = 20000
    off, chars, _, _ = datagen.strgen(dkeys, seed=21, shuffle_seed=2)
    rng = np.random.default_rng(6)
    flat = datagen.take_rows(off, chars, rng.permutation(np.repeat(np.arange(dkeys), 2)))
    zipf = datagen.take_rows(off, chars, datagen.zipf_indices(150000, dkeys, 1.0, seed=8))
    (roff, rchars), (soff, schars) = (flat, zipf) if skewed_side == "probe" else (zipf, flat)
    out_r, out_s = gpu_join_strings(ctx, roff, rchars, soff, schars)
    assert ctx.timings()["bucket_join"] == 0, "the bucket join cannot hold a Zipf relation"
    t_r, t_s, _ = oracle.full_join(Keys(offsets=roff, chars=rchars), Keys(offsets=soff, chars=schars))
    assert out_r.size == 300000
    c_r, c_s = oracle.sort_pairs(out_r, out_s)
    e_r, e_s = oracle.sort_pairs(t_r, t_s)
    assert np.array_equal(c_r, e_r) and np.array_equal(c_s, e_s)
    # same with u64 keys (identity hash)
    ku = datagen.unique_u64(dkeys, seed=3)
    a = ku[rng.permutation(np.repeat(np.arange(dkeys), 2))]
    b = ku[datagen.zipf_indices(150000, dkeys, 1.0, seed=9)]
    check_operator_u64(ctx, oracle, *((a, b) if skewed_side == "probe" else (b, a)))


def test_join_golden_duplicates(ctx, oracle):
    g = G.load("dup_join_u64.json")
    rk, sk = G.u64(g["r_keys"]), G.u64(g["s_keys"])
    out_r, out_s = check_join_u64(ctx, oracle, rk, sk)
    # every pair the reference emits is in our result (ours is the full cross product)
    assert set(zip(g["join_r"], g["join_s"])) <= set(zip(out_r.tolist(), out_s.tolist()))


@pytest.mark.parametrize("nr,ns,rng_hi", [(0, 10, 5), (10, 0, 5), (1, 1, 1), (1, 3, 1), (5000, 7000, 600),
                                          (30000, 100, 50), (100, 30000, 50), (20000, 20000, 2 ** 63)])
def test_join_shapes(ctx, oracle, nr, ns, rng_hi):
    rng = np.random.default_rng(nr * 7 + ns)
    check_join_u64(ctx, oracle, rng.integers(0, rng_hi, size=nr, dtype=np.uint64),
                   rng.integers(0, rng_hi, size=ns, dtype=np.uint64))


def test_join_runs_longer_than_a_tile(ctx, oracle):
    # a 9000-row S run against 3 R rows and vice versa: runs cross many merge-path tiles
    rk = np.concatenate([np.full(3, 42), np.arange(100, 6000), np.full(9000, 77)]).astype(np.uint64)
    sk = np.concatenate([np.full(9000, 42), np.arange(3000, 9000), np.full(2, 77)]).astype(np.uint64)
    check_join_u64(ctx, oracle, rk, sk)


def test_join_where_reference_is_exact(ctx, oracle):
    # unique keys + maximal non-matching sentinels: here oracle walk == reference == true join
    keys = datagen.unique_u64(20000, seed=8) >> np.uint64(1)
    rk = np.concatenate([datagen.permuted(keys, 3), [np.uint64(2 ** 64 - 1)]]).astype(np.uint64)
    sk = np.concatenate([datagen.permuted(keys, 4), [np.uint64(2 ** 64 - 2)]]).astype(np.uint64)
    out_r, out_s = check_join_u64(ctx, oracle, rk, sk)
    r, s = Keys(u64=rk), Keys(u64=sk)
    rh, rr = oracle.radix_sort(rk, r)
    sh, sr = oracle.radix_sort(sk, s)
    w_r, w_s = oracle.merge_walk_reference(rh, rr, sh, sr, r, s)
    assert all(np.array_equal(a, b) for a, b in zip(oracle.sort_pairs(w_r, w_s), oracle.sort_pairs(out_r, out_s)))


def gpu_join_strings(ctx, off, chars, off2, chars2):
    d_off, d_chars, d_off2, d_chars2 = dev(off), dev(chars), dev(off2), dev(chars2)
    m = ctx.join_strings(d_off, d_chars, d_off2, d_chars2)
    return ctx.join_result_to_host(m)


def test_gather_strings_puts_keys_in_row_order(ctx):
    off, chars, _, _ = datagen.strgen(70000, seed=19, shuffle_seed=4)
    rng = np.random.default_rng(2)
    for rows in (rng.integers(0, 70000, size=100001), np.arange(70000)[::-1].copy(), np.empty(0, np.int64)):
        g_off, g_chars = ctx.gather_strings(dev(off), dev(chars), dev(rows.astype(np.uint32)))
        e_off, e_chars = datagen.take_rows(off, chars, rows)
        assert np.array_equal(g_off.cpu().numpy().view(np.uint64), e_off)
        assert np.array_equal(g_chars.cpu().numpy(), e_chars)


def test_join_strings_of_one_hash_range(ctx, oracle):
    # the local step of the multi-GPU string join: every key's std::hash has the same top 2 bits
    off, chars, _, _ = datagen.strgen(400000, seed=23, shuffle_seed=5)
    h = oracle.hash_strings(off, chars)
    mine = np.flatnonzero((h >> np.uint64(62)) == 1)
    rng = np.random.default_rng(8)
    roff, rchars = datagen.take_rows(off, chars, rng.permutation(np.concatenate([mine, mine[:5000]])))
    soff, schars = datagen.take_rows(off, chars, rng.permutation(mine)[: mine.size * 3 // 4])
    t_r, t_s, _ = oracle.full_join(Keys(offsets=roff, chars=rchars), Keys(offsets=soff, chars=schars))
    e_r, e_s = oracle.sort_pairs(t_r, t_s)
    for bits, uniform in ((2, True), (0, False)):
        m = ctx.join_strings(dev(roff), dev(rchars), dev(soff), dev(schars), prefix_bits=bits)
        assert (ctx.timings()["bucket_join"] == 1) == uniform
        c_r, c_s = oracle.sort_pairs(*ctx.join_result_to_host(m))
        assert np.array_equal(c_r, e_r) and np.array_equal(c_s, e_s)
        # the device-side copy of the result is the same list
        d_r, d_s = ctx.join_result_tensors(m)
        c_r, c_s = oracle.sort_pairs(d_r.cpu().numpy().view(np.uint32), d_s.cpu().numpy().view(np.uint32))
        assert np.array_equal(c_r, e_r) and np.array_equal(c_s, e_s)


def test_join_strings_golden(ctx, oracle):
    g = G.load("strgen_join_600.json")
    off, chars = np.array(g["r_offsets"], np.uint64), G.u8(g["r_chars_hex"])
    off2, chars2 = np.array(g["s_offsets"], np.uint64), G.u8(g["s_chars_hex"])
    out_r, out_s = gpu_join_strings(ctx, off, chars, off2, chars2)
    ours = set(zip(out_r.tolist(), out_s.tolist()))
    assert set(zip(g["join_r"], g["join_s"])) <= ours and len(ours) == 600


def test_join_strings_vs_oracle_with_duplicates_and_hash_collisions(ctx, oracle):
    off, chars, _, _ = datagen.strgen(50000, seed=6, shuffle_seed=1)
    rng = np.random.default_rng(3)
    pick_r = rng.integers(0, 50000, size=60000)   # duplicates on both sides (N:M)
    pick_s = rng.integers(0, 50000, size=40000)
    roff, rchars = datagen.take_rows(off, chars, pick_r)
    soff, schars = datagen.take_rows(off, chars, pick_s)
    out_r, out_s = gpu_join_strings(ctx, roff, rchars, soff, schars)
    t_r, t_s, _ = oracle.full_join(Keys(offsets=roff, chars=rchars), Keys(offsets=soff, chars=schars))
    c_r, c_s = oracle.sort_pairs(out_r, out_s)
    assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)


def test_join_verifies_keys_on_forced_hash_collisions(ctx, oracle):
    # same hash column, different keys: feed the merge join pre-"hashed" columns that collide
    items_r = [b"alpha", b"beta", b"gamma", b"delta"]
    items_s = [b"alpha", b"BETA", b"gamma", b"delta", b"alpha"]
    roff, rchars = datagen.strings_from_list(items_r)
    soff, schars = datagen.strings_from_list(items_s)
    rh = np.array([5, 5, 5, 9], np.uint64)       # all of alpha/beta/gamma "collide"
    sh = np.array([5, 5, 5, 9, 5], np.uint64)
    d = lambda a: dev(np.concatenate([a, np.zeros(8, a.dtype)])) if a.dtype == np.uint8 else dev(a)
    rhs, rrs = ctx.sort_pairs(dev(rh))
    shs, srs = ctx.sort_pairs(dev(sh))
    R = hmj.SortedRelation(rhs, rrs, hmj.KEY_STRING, d(roff), d(rchars))
    S = hmj.SortedRelation(shs, srs, hmj.KEY_STRING, d(soff), d(schars))
    out_r, out_s = ctx.merge_join(R, S)
    got = sorted(zip(host(out_r, np.uint32).tolist(), host(out_s, np.uint32).tolist()))
    assert got == [(0, 0), (0, 4), (2, 2), (3, 3)]
    assert ctx.merge_join_count(R, S) == 4


# ------------------------------------------------------------------ whole operator, host buffers
def test_join_host_api(ctx, oracle):
    keys = datagen.unique_u64(50000, seed=2)
    rk, sk = datagen.permuted(keys, 5), datagen.permuted(keys, 6)[:40000]
    out_r, out_s = np.empty(50000, np.uint32), np.empty(50000, np.uint32)
    m = ctx.join_u64_host(rk, sk, out_r, out_s)
    assert m == 40000
    assert np.array_equal(rk[out_r[:m]], sk[out_s[:m]])
    t_r, t_s, _ = oracle.full_join(Keys(u64=rk), Keys(u64=sk))
    c_r, c_s = oracle.sort_pairs(out_r[:m], out_s[:m])
    assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)
    with pytest.raises(hmj.HmjError):
        ctx.join_u64_host(rk, sk, np.empty(10, np.uint32), np.empty(10, np.uint32))  # HMJ_ECAPACITY
    off, chars, _, _ = datagen.strgen(30000, seed=7, shuffle_seed=1)
    off2, chars2, _, _ = datagen.strgen(30000, seed=7, shuffle_seed=2)
    out_r, out_s = np.empty(30000, np.uint32), np.empty(30000, np.uint32)
    assert ctx.join_strings_host(off, chars, off2, chars2, out_r, out_s) == 30000
    t_r, t_s, _ = oracle.full_join(Keys(offsets=off, chars=chars), Keys(offsets=off2, chars=chars2))
    c_r, c_s = oracle.sort_pairs(out_r, out_s)
    assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)


def test_host_joins_in_flight_from_two_threads(oracle):
    # two host threads, a context (stream + workspace) each, host-buffer joins at the same time:
    # the copy turns of host_api.cu must not mix up anyone's buffers
    import threading
    jobs = []
    for seed in (3, 4):
        keys = datagen.unique_u64(300000, seed=seed)
        rk, sk = datagen.permuted(keys, 1).copy(), datagen.permuted(keys, 2)[:250000].copy()
        rk[:100] = rk[100:200]
        jobs.append((rk, sk, np.empty(300000, np.uint32), np.empty(300000, np.uint32)))
    ctxs = [hmj.Context(0, use_torch_stream=False) for _ in jobs]
    got = [[] for _ in jobs]
    errors = []

    def lane(i):
        try:
            rk, sk, out_r, out_s = jobs[i]
            for _ in range(6):
                got[i].append(ctxs[i].join_u64_host(rk, sk, out_r, out_s))
        except Exception as e:  # surfaced below: an assert in a thread would go unseen
            errors.append(e)

    threads = [threading.Thread(target=lane, args=(i,)) for i in range(len(jobs))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for i, (rk, sk, out_r, out_s) in enumerate(jobs):
        t_r, t_s, _ = oracle.full_join(Keys(u64=rk), Keys(u64=sk))
        assert got[i] == [t_r.size] * 6
        c_r, c_s = oracle.sort_pairs(out_r[:t_r.size], out_s[:t_r.size])
        assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)
        ctxs[i].close()


def test_sort_host_api(ctx, oracle):
    keys = np.random.default_rng(8).integers(0, 2 ** 64, size=77777, dtype=np.uint64)
    h, r = ctx.hash_sort_u64_host(keys)
    assert np.array_equal(h, np.sort(keys)) and np.array_equal(keys[r], h)
    h2, r2 = ctx.sort_pairs_host(keys, np.arange(keys.size, dtype=np.uint32)[::-1].copy())
    assert np.array_equal(h2, h)


# ------------------------------------------------------------------ A6 partitioning
def test_partition_reference_case_and_oracle(ctx, oracle):
    top = np.uint64(1) << np.uint64(63)   # partitioned_hash_test.cc:14-46
    i = np.arange(5, dtype=np.uint64)
    keys = np.stack([i, i | top, i | top | np.uint64(1024)], axis=1).reshape(-1)
    assert ctx.partition_counts_host(keys, 1).tolist() == [5, 10]
    g = G.load("partition_u64.json")
    keys = G.u64(g["keys"])
    for bits, counts in g["counts"].items():
        assert ctx.partition_counts_host(keys, int(bits)).tolist() == counts
    big = np.random.default_rng(1).integers(0, 2 ** 64, size=300000, dtype=np.uint64)
    for bits in (1, 3, 8, 12, 14, 16):
        assert np.array_equal(ctx.partition_counts_host(big, bits), oracle.partition_counts(big, bits))
    for bits in (1, 3, 8):
        oh, orow, offs = ctx.partition_pairs(dev(big), None, bits)
        oh, orow, offs = host(oh, np.uint64), host(orow, np.uint32), host(offs, np.uint64)
        part = (big >> np.uint64(64 - bits)).astype(np.int64)
        order = np.argsort(part, kind="stable")
        assert np.array_equal(orow, order.astype(np.uint32)) and np.array_equal(oh, big[order])
        assert np.array_equal(np.diff(offs), oracle.partition_counts(big, bits))


# ------------------------------------------------------------------ A7 hash-join baseline
def test_hashjoin_baseline_vs_oracle(ctx, oracle):
    rng = np.random.default_rng(12)
    rk = rng.integers(0, 30000, size=50000, dtype=np.uint64)
    sk = rng.integers(0, 30000, size=20000, dtype=np.uint64)
    sk[5] = np.uint64(2 ** 64 - 1)
    rk[7] = np.uint64(2 ** 64 - 1)
    R = hmj.SortedRelation(dev(rk), None)
    S = hmj.SortedRelation(dev(sk), None)
    assert ctx.hashjoin_build_probe(R, S) == oracle.hash_join_raw(Keys(u64=rk), Keys(u64=sk))
    off, chars, _, _ = datagen.strgen(20000, seed=6)
    roff, rchars = datagen.take_rows(off, chars, rng.integers(0, 20000, size=30000))
    soff, schars = datagen.take_rows(off, chars, rng.integers(0, 20000, size=15000))
    rkeys, skeys = Keys(offsets=roff, chars=rchars), Keys(offsets=soff, chars=schars)
    R = hmj.SortedRelation(dev(oracle.hashes(rkeys)), None, hmj.KEY_STRING, dev(roff), dev(rchars))
    S = hmj.SortedRelation(dev(oracle.hashes(skeys)), None, hmj.KEY_STRING, dev(soff), dev(schars))
    assert ctx.hashjoin_build_probe(R, S) == oracle.hash_join_raw(rkeys, skeys)


# ------------------------------------------------------------------ device generators (8 f4)
def test_device_generators_match_their_numpy_restatements(ctx):
    for domain, seed in ((1, 3), (2, 9), (1000, 1), (4097, 77), (1 << 20, 123456), (3 * 10 ** 6 + 11, 5)):
        n = min(domain, 300000)
        got = host(ctx.gen_permute(n, domain, seed, base=domain - n), np.uint32)
        want = datagen.feistel_permute(np.arange(domain - n, domain, dtype=np.uint32), domain, seed)
        assert np.array_equal(got, want)
        if n == domain:
            assert np.array_equal(np.sort(got), np.arange(domain, dtype=np.uint32)), "not a bijection"
    src = np.random.default_rng(2).integers(0, 5000, size=70000, dtype=np.uint32)
    assert np.array_equal(host(ctx.gen_permute(src.size, 5000, 11, src=dev(src)), np.uint32),
                          datagen.feistel_permute(src, 5000, 11))
    cdf = datagen.zipf_cdf(30000, 1.0)
    d_cdf = torch.from_numpy(cdf).cuda()
    got = host(ctx.gen_zipf(d_cdf, 200000, seed=7, row_base=12345), np.uint32)
    assert np.array_equal(got, datagen.zipf_ranks(cdf, 200000, 7, 12345))
    assert 0.07 < (got == 0).mean() < 0.12    # head key of Zipf(1.0) over 30000 keys ~ 9.2 %
    # strgen keys: base words and w_i-w_j pairs, any id order, duplicates allowed
    gen = datagen.DeviceStrgen(ctx, 10000, seed=21)
    ids = np.concatenate([np.arange(0, 230), np.random.default_rng(3).integers(0, 10000, size=3000)]).astype(np.uint32)
    off, chars = gen.keys(dev(ids))
    w_off, w_chars = datagen.strgen_keys(gen.h_woff, gen.h_wchars, ids)
    assert np.array_equal(host(off, np.uint64), w_off) and np.array_equal(chars.cpu().numpy(), w_chars)
    # ... and the unshuffled relation is exactly the host generator's
    off, chars = gen.keys(dev(np.arange(10000, dtype=np.uint32)))
    h_off, h_chars, _ = datagen.strgen_unshuffled(10000, seed=21)
    assert np.array_equal(host(off, np.uint64), h_off) and np.array_equal(chars.cpu().numpy(), h_chars)


def test_device_generated_string_join_against_oracle(ctx, oracle):
    # C5's shape from the device generators: every key twice on the build side, Zipf(1.0) probe
    gen = datagen.DeviceStrgen(ctx, 30000, seed=5)
    r_ids = gen.shuffled_ids(60000, seed=1, copies=2)
    cdf = torch.from_numpy(datagen.zipf_cdf(30000, 1.0)).cuda()
    s_ids = gen.zipf_ids(cdf, 60000, seed=2, key_seed=3)
    assert np.array_equal(np.bincount(host(r_ids, np.uint32), minlength=30000), np.full(30000, 2))
    (roff, rchars), (soff, schars) = gen.keys(r_ids), gen.keys(s_ids)
    m = ctx.join_strings(roff, rchars, soff, schars)
    assert m == 2 * 60000
    c_r, c_s = oracle.sort_pairs(*ctx.join_result_to_host(m))
    t_r, t_s, _ = oracle.full_join(Keys(offsets=host(roff, np.uint64), chars=rchars.cpu().numpy()),
                                   Keys(offsets=host(soff, np.uint64), chars=schars.cpu().numpy()))
    e_r, e_s = oracle.sort_pairs(t_r, t_s)
    assert np.array_equal(c_r, e_r) and np.array_equal(c_s, e_s)


# ------------------------------------------------------------------ benchmark-sized parity
def test_join_u64_64m_rows_against_the_compiled_reference(ctx, ref):
    # 64M x 64M unique keys: the 8+8-bit two-level plan with 16384 buckets that the benchmark's
    # shape uses, against the UNMODIFIED reference (exact here: unique keys, hashjoin.h:70-180)
    if ref is None:
        pytest.skip("oracle/_ref not built")
    n = 1 << 26
    keys = datagen.unique_u64(n, seed=29)
    rk, sk = datagen.permuted(keys, 1), datagen.permuted(keys, 2)
    m = ctx.join_u64(dev(rk), dev(sk))
    tm = ctx.timings()
    assert tm["bucket_join"] == 1 and tm["scatter_levels"] == 4
    out_r, out_s = ctx.join_result_to_host(m)
    ctx.lib.hmj_release_workspace(ctx.h)
    e_r, e_s, _, _ = ref.join(Keys(u64=rk), Keys(u64=sk), threads=ref.cores())
    # the reference walks in hash order and stops at the last element of either side (8a A5):
    # with unique keys it may miss the very last match only
    assert m == n and e_r.size >= n - 1
    order = np.argsort(out_r, kind="stable")
    g_s = np.empty(n, np.uint32)
    g_s[out_r[order]] = out_s[order]          # probe row of every build row (keys are unique)
    assert np.array_equal(np.sort(out_r), np.arange(n, dtype=np.uint32))
    assert np.array_equal(g_s[e_r], e_s), "pairs differ from the reference's"
    assert np.array_equal(rk[out_r], sk[out_s])


def test_join_strings_8m_rows_against_the_compiled_reference(ctx, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built")
    n = 1 << 23
    gen = datagen.DeviceStrgen(ctx, n, seed=9)
    (roff, rchars), (soff, schars) = gen.keys(gen.shuffled_ids(n, seed=1)), gen.keys(gen.shuffled_ids(n, seed=2))
    m = ctx.join_strings(roff, rchars, soff, schars)
    assert ctx.timings()["bucket_join"] == 1
    out_r, out_s = ctx.join_result_to_host(m)
    R = Keys(offsets=host(roff, np.uint64), chars=rchars.cpu().numpy())
    S = Keys(offsets=host(soff, np.uint64), chars=schars.cpu().numpy())
    ctx.lib.hmj_release_workspace(ctx.h)
    e_r, e_s, _, _ = ref.join(R, S, threads=ref.cores())
    assert m == n and e_r.size >= n - 1
    g_s = np.empty(n, np.uint32)
    g_s[out_r] = out_s
    assert np.array_equal(np.sort(out_r), np.arange(n, dtype=np.uint32))
    assert np.array_equal(g_s[e_r], e_s), "pairs differ from the reference's"


def test_sort_after_release_workspace_with_scribbled_memory(ctx, oracle):
    # the look-back descriptors of the stable passes carry epoch tags instead of being cleared per
    # pass: a released workspace must not leave a stale "seen" state behind
    rng = np.random.default_rng(14)
    keys = rng.integers(0, 1 << 20, size=600000, dtype=np.uint64)   # skewed: takes the stable passes
    h0, r0 = ctx.sort_pairs(dev(keys))
    assert ctx.timings()["sort_passes"] > 0
    want_h, want_r = host(h0, np.uint64).copy(), host(r0, np.uint32).copy()
    del h0, r0
    for word in (0xFFFFFFFF, 0x00010001, 0x7FFF7FFF):
        ctx.lib.hmj_release_workspace(ctx.h)
        torch.cuda.empty_cache()
        junk = torch.full((96 << 20,), word - (1 << 32) if word >= 1 << 31 else word, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        del junk
        torch.cuda.empty_cache()
        h, r = ctx.sort_pairs(dev(keys))
        assert np.array_equal(host(h, np.uint64), want_h) and np.array_equal(host(r, np.uint32), want_r)


# ------------------------------------------------------------------ plan selection by sampling
def _check_pairs_1to1(rk, sk, out_r, out_s):
    assert out_r.size == rk.size
    assert np.array_equal(rk[out_r], sk[out_s])
    assert np.array_equal(np.sort(out_r), np.arange(rk.size, dtype=np.uint32))
    assert np.array_equal(np.sort(out_s), np.arange(sk.size, dtype=np.uint32))


def test_join_dense_integer_keys_take_the_bucket_path(ctx):
    # identity hash of keys 0..N-1: the upper 64 - log2(N) hash bits are all zero.  The sample finds
    # the shared prefix, the plan skips it, level 1 checks it on every row.
    n = 1 << 23
    rng = np.random.default_rng(3)
    rk, sk = rng.permutation(n).astype(np.uint64), rng.permutation(n).astype(np.uint64)
    m = ctx.join_u64(dev(rk), dev(sk))
    tm = ctx.timings()
    assert tm["bucket_join"] == 1 and tm["sort_passes"] == 0, "dense keys did not take the bucket path"
    _check_pairs_1to1(rk, sk, *ctx.join_result_to_host(m))
    # the same keys behind a constant offset that leaves a common prefix
    off = np.uint64(0xABCD) << np.uint64(40)
    m = ctx.join_u64(dev(rk + off), dev(sk + off))
    assert ctx.timings()["bucket_join"] == 1
    _check_pairs_1to1(rk, sk, *ctx.join_result_to_host(m))
    # one row the sample does not see breaks the prefix: the attempt is void, the result is not
    rk2 = rk.copy()
    rk2[12345] = np.uint64(1) << np.uint64(63)
    m = ctx.join_u64(dev(rk2), dev(sk))
    assert ctx.timings()["bucket_join"] == 0
    out_r, out_s = ctx.join_result_to_host(m)
    assert m == n - 1 and np.array_equal(rk2[out_r], sk[out_s])
    assert np.unique(out_r).size == n - 1


def test_join_dense_keys_of_any_count_take_the_bucket_path(ctx):
    # ids 0..N-1 with N not a power of two fill only N / 2^k of the range under their prefix: the
    # sample measures that density and the plan provides for it (bucket count, level-1 regions)
    rng = np.random.default_rng(31)
    for n, base in ((5_000_000, 0), (3 << 21, 1 << 40), ((1 << 22) + 1000, 0), (4_500_000, (1 << 23) - 2_000_000)):
        rk = rng.permutation(n).astype(np.uint64) + np.uint64(base)
        sk = rng.permutation(n).astype(np.uint64) + np.uint64(base)
        m = ctx.join_u64(dev(rk), dev(sk))
        tm = ctx.timings()
        assert tm["bucket_join"] == 1 and tm["sort_passes"] == 0 and tm["sort_fallbacks"] == 0, (n, base, tm)
        _check_pairs_1to1(rk, sk, *ctx.join_result_to_host(m))
    # a probe side that is a third of the ids, spread over the range
    n = 6_000_000
    rk = rng.permutation(n).astype(np.uint64)
    sk = rng.choice(n, size=n // 3, replace=False).astype(np.uint64)
    m = ctx.join_u64(dev(rk), dev(sk))
    assert m == sk.size and ctx.timings()["bucket_join"] == 1
    out_r, out_s = ctx.join_result_to_host(m)
    assert np.array_equal(rk[out_r], sk[out_s]) and np.unique(out_s).size == m
    # through the host entry (chunks of R): 3M dense ids
    n = 3 << 20
    rk, sk = rng.permutation(n).astype(np.uint64), rng.permutation(n).astype(np.uint64)
    o_r, o_s = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    assert ctx.join_u64_host(rk, sk, o_r, o_s) == n
    assert ctx.timings()["bucket_join"] == 1 and ctx.timings()["scatter_levels"] > 4, "dense keys left the chunked plan"
    _check_pairs_1to1(rk, sk, o_r, o_s)


def test_join_skewed_keys_skip_the_scatter_attempt(ctx, oracle):
    # a Zipf-distributed integer relation: the sample sees the head key's region at several times
    # its share, so no bucket scatter is attempted at all
    n = 1 << 22
    keys = datagen.unique_u64(1 << 16, seed=4)
    rk = keys[datagen.zipf_ranks(datagen.zipf_cdf(keys.size), n, 3).astype(np.int64)]
    sk = datagen.permuted(np.concatenate([keys] * 2), 5)
    m = ctx.join_u64(dev(rk), dev(sk))
    tm = ctx.timings()
    # (the uniform relation S is still sorted by its own two scatter levels; R goes straight to the
    # histogram-planned stable passes)
    assert tm["bucket_join"] == 0 and tm["scatter_levels"] <= 2 and tm["sort_fallbacks"] == 0, \
        "a scatter attempt was thrown away on a relation the sample had found skewed"
    assert m == 2 * n
    out_r, out_s = ctx.join_result_to_host(m)
    assert np.array_equal(rk[out_r], sk[out_s])
    w = (out_r.astype(np.uint64) << np.uint64(32)) | out_s.astype(np.uint64)
    assert np.unique(w).size == m


def test_join_order_option_gives_hash_order_reproducibly(ctx, oracle):
    n = 300000
    keys = datagen.unique_u64(n, seed=8)
    rk, sk = datagen.permuted(keys, 1).copy(), datagen.permuted(keys, 2).copy()
    sk[:5000] = sk[5000:10000]   # duplicates on the probe side
    runs = []
    ctx.set_join_order(True)
    try:
        for _ in range(2):
            m = ctx.join_u64(dev(rk), dev(sk))
            assert ctx.timings()["bucket_join"] == 0
            runs.append(ctx.join_result_to_host(m))
    finally:
        ctx.set_join_order(False)
    (a_r, a_s), (b_r, b_s) = runs
    assert np.array_equal(a_r, b_r) and np.array_equal(a_s, b_s), "hash order is not reproducible"
    h = rk[a_r]
    assert np.all(h[1:] >= h[:-1]), "pairs are not in ascending hash order"
    t_r, t_s, _ = oracle.full_join(Keys(u64=rk), Keys(u64=sk))
    c_r, c_s = oracle.sort_pairs(a_r, a_s)
    assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)
    m = ctx.join_u64(dev(rk), dev(sk))
    assert ctx.timings()["bucket_join"] == 1, "the default order must take the faster plan again"


def test_host_join_in_chunks_and_its_fallbacks(ctx, oracle):
    # hmj_join_u64_host copies R in chunks, joins every chunk against S's buckets and sends the
    # pairs home while the next chunk arrives; the union is the join
    n = 3 << 20
    keys = datagen.unique_u64(n, seed=12)
    rk, sk = datagen.permuted(keys, 1).copy(), datagen.permuted(keys, 2).copy()
    out_r, out_s = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    assert ctx.join_u64_host(rk, sk, out_r, out_s) == n
    assert ctx.timings()["bucket_join"] == 1 and ctx.timings()["scatter_levels"] > 4, "not the chunked plan"
    _check_pairs_1to1(rk, sk, out_r, out_s)
    # R much smaller than S and the other way round
    for a, b in ((rk[: n // 5], sk), (rk, sk[: n // 7])):
        o_r, o_s = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        m = ctx.join_u64_host(a, b, o_r, o_s)
        assert m == min(a.size, b.size) and np.array_equal(a[o_r[:m]], b[o_s[:m]])
        assert np.unique(o_r[:m]).size == m
    # every key three times on the probe side: 3n pairs do not fit the first guess, the plain path
    # re-runs with the exact size
    sk3 = datagen.permuted(np.concatenate([sk, sk, sk]), 9)
    o_r, o_s = np.zeros(3 * n, np.uint32), np.zeros(3 * n, np.uint32)
    assert ctx.join_u64_host(rk, sk3, o_r, o_s) == 3 * n
    assert np.array_equal(rk[o_r], sk3[o_s])
    w = (o_r.astype(np.uint64) << np.uint64(32)) | o_s.astype(np.uint64)
    assert np.unique(w).size == 3 * n
    # dense keys 0..N-1 through the host entry: the sample of S and of R's first chunk finds the
    # shared prefix, so the chunked plan (not the copy-join-copy fallback) runs
    # (a power-of-two range here; other counts: test_join_dense_keys_of_any_count_take_the_bucket_path)
    rng = np.random.default_rng(21)
    nd = 1 << 22
    dk_r, dk_s = rng.permutation(nd).astype(np.uint64), rng.permutation(nd).astype(np.uint64)
    d_r, d_s = np.zeros(nd, np.uint32), np.zeros(nd, np.uint32)
    assert ctx.join_u64_host(dk_r, dk_s, d_r, d_s) == nd
    assert ctx.timings()["bucket_join"] == 1 and ctx.timings()["scatter_levels"] > 4, "dense keys left the chunked plan"
    _check_pairs_1to1(dk_r, dk_s, d_r, d_s)
    # a capacity that is too small is reported, with the number of pairs
    with pytest.raises(hmj.HmjError) as ei:
        ctx.join_u64_host(rk, sk, out_r[: n // 2], out_s[: n // 2])
    assert ei.value.code == -5
