"""Pins the CPU oracle (oracle/hmj_oracle.c) against the UNMODIFIED reference compiled into
oracle/_ref/libhmj_ref.so: the reference's own test cases (radix_hash_test.cc,
radix_sort_test.cc, partitioned_hash_test.cc) plus seeded random inputs.  CPU only."""
import numpy as np
import pytest

from hashmergejoin_b200 import datagen
from tests.oracle_binding import Keys

pytestmark = pytest.mark.skipif(False, reason="")


@pytest.fixture(autouse=True)
def _need_ref(ref):
    if ref is None:
        pytest.skip("oracle/_ref not built (reference not mounted)")


SURVEY_GOLDEN = {  # SURVEY.md section 8(c), std::hash<std::string> on libstdc++ x86-64
    b"": 0x553e93901e462a6e, b"a": 0x454ddee488c1ed6b, b"abc": 0x32d82bf8ed3dba39,
    b"hello-world": 0x4266939658a8b675, b"aardvark-zebra": 0xbdf84b350391737a,
    b"0123456789abcdef": 0x934b0d760642ed86, b"the quick brown fox": 0x265845ddfb0280cd,
}


def test_std_hash_golden(oracle, ref):
    for s, h in SURVEY_GOLDEN.items():
        assert ref.hash_bytes(s) == h
        assert oracle.hash_bytes(s) == h


def test_std_hash_random_all_lengths(oracle, ref):
    rng = np.random.default_rng(11)
    items = [bytes(rng.integers(0, 256, size=ln, dtype=np.uint8)) for ln in range(0, 130) for _ in range(8)]
    off, chars = datagen.strings_from_list(items)
    assert np.array_equal(oracle.hash_strings(off, chars), ref.hash_strings(off, chars))


def test_std_hash_strgen(oracle, ref):
    off, chars, _, _ = datagen.strgen(20000, seed=5)
    assert np.array_equal(oracle.hash_strings(off, chars), ref.hash_strings(off, chars))


def test_optimal_partition(oracle, ref):
    ns = list(range(0, 300)) + [10 ** k for k in range(3, 10)] + [2 ** k for k in range(6, 34)] + \
        [2 ** k - 1 for k in range(6, 34)] + [1000003, 64 * 10 ** 6, 256 * 2 ** 20, 512 * 2 ** 20]
    for n in ns:
        assert oracle.optimal_partition(n) == ref.lib.ref_optimal_partition(n), n


# ---- radix_hash_test.cc cases (identity hash on ints == std::hash<uint64_t>)
REF_SORT_CASES = [
    (np.arange(4, -1, -1), 3), (np.arange(5, 0, -1), 1), (np.arange(5, 0, -1), 2),
    (np.arange(12345, 0, -1), 1), (np.arange(12345, 0, -1), 14), (np.arange(1 << 18, 0, -1), 0),
]


@pytest.mark.parametrize("case", range(len(REF_SORT_CASES)))
@pytest.mark.parametrize("threads", [1, 4])
def test_non_inplace_reference_cases(oracle, ref, case, threads):
    keys, bits = REF_SORT_CASES[case]
    k = Keys(u64=keys.astype(np.uint64))
    h_ref, r_ref, _ = ref.sort(k, threads=threads, bits=bits)
    assert np.array_equal(h_ref, np.sort(k.u64))  # what the reference test asserts
    h_or, r_or = oracle.radix_sort(k.u64, k, bits=bits)
    assert np.array_equal(h_or, h_ref)
    assert np.array_equal(r_or, r_ref)  # exact element order, not only the hash column


@pytest.mark.parametrize("bits", [0, 1, 6, 8, 11])
@pytest.mark.parametrize("n", [1, 2, 65, 1 << 16])
def test_non_inplace_random_u64(oracle, ref, n, bits):
    rng = np.random.default_rng(n * 31 + bits)
    k = Keys(u64=rng.integers(0, 2 ** 64, size=n, dtype=np.uint64))
    h_ref, r_ref, _ = ref.sort(k, threads=3, bits=bits)
    h_or, r_or = oracle.radix_sort(k.u64, k, bits=bits)
    assert np.array_equal(h_or, h_ref) and np.array_equal(r_or, r_ref)


def test_non_inplace_duplicates_small_range(oracle, ref):
    rng = np.random.default_rng(3)
    k = Keys(u64=rng.integers(0, 50, size=5000, dtype=np.uint64))
    h_ref, r_ref, _ = ref.sort(k, threads=2, bits=0)
    h_or, r_or = oracle.radix_sort(k.u64, k, bits=0)
    assert np.array_equal(h_or, h_ref) and np.array_equal(r_or, r_ref)


def test_non_inplace_strings(oracle, ref):
    off, chars, _, _ = datagen.strgen(30000, seed=9)
    k = Keys(offsets=off, chars=chars)
    h_ref, r_ref, _ = ref.sort(k, threads=4, bits=0)
    h_or, r_or = oracle.radix_sort(oracle.hash_strings(off, chars), k, bits=0)
    assert np.array_equal(h_or, h_ref) and np.array_equal(r_or, r_ref)


@pytest.mark.parametrize("case", range(len(REF_SORT_CASES)))
def test_inplace_seq_reference_cases(oracle, ref, case):
    keys, bits = REF_SORT_CASES[case]
    k = Keys(u64=keys.astype(np.uint64))
    h_ref, r_ref, _ = ref.sort_inplace(k.u64, mode=0, bits=bits)
    assert np.array_equal(h_ref, np.sort(k.u64))
    h_or, r_or = oracle.radix_sort(k.u64, k, bits=bits, inplace_seq=True)
    assert np.array_equal(h_or, h_ref) and np.array_equal(r_or, r_ref)


def test_inplace_seq_random(oracle, ref):
    rng = np.random.default_rng(5)
    k = Keys(u64=rng.integers(0, 2 ** 64, size=1 << 16, dtype=np.uint64))
    h_ref, r_ref, _ = ref.sort_inplace(k.u64, mode=0)
    h_or, r_or = oracle.radix_sort(k.u64, k, inplace_seq=True)
    assert np.array_equal(h_or, h_ref) and np.array_equal(r_or, r_ref)


def test_inplace_par_reference_cases(oracle, ref):
    # radix_inplace_par_test: 2^18 reversed / 4 threads; top-bit input / 8 threads / 1 bit;
    # 2^16 random / all cores / 1 bit.  The parallel permutation order is nondeterministic, so
    # only the hash column (what the reference asserts) and the pair multiset are compared.
    rng = np.random.default_rng(6)
    top = np.empty(1024, np.uint64)
    top[0::2] = (np.uint64(1) << np.uint64(63)) + np.arange(512, dtype=np.uint64)
    top[1::2] = np.arange(512, dtype=np.uint64)
    cases = [(np.arange(1 << 18, 0, -1).astype(np.uint64), 4, 0), (top, 8, 1),
             (rng.integers(0, 2 ** 64, size=1 << 16, dtype=np.uint64), ref.cores(), 1)]
    for keys, threads, bits in cases:
        h_ref, r_ref, _ = ref.sort_inplace(keys, mode=1, threads=threads, bits=bits)
        assert np.array_equal(h_ref, np.sort(keys))
        h_or, r_or = oracle.radix_sort(keys, Keys(u64=keys), bits=bits)
        assert np.array_equal(h_or, h_ref)
        assert all(np.array_equal(a, b) for a, b in zip(oracle.canonical(h_or, r_or), oracle.canonical(h_ref, r_ref)))


def test_radix_sort_h_reference_cases(ref):
    # radix_sort_test.cc: 2^16 random pairs, in-place and non-in-place, key column vs std::sort
    rng = np.random.default_rng(7)
    keys = rng.integers(0, 2 ** 64, size=1 << 16, dtype=np.uint64)
    vals = np.arange(keys.size, dtype=np.uint32)
    for mode in (0, 1):
        k, v, _ = ref.sort_int_pairs(keys, vals, mode, threads=ref.cores())
        assert np.array_equal(k, np.sort(keys))
        assert np.array_equal(keys[v], k)


# ---- partitioned_hash_test.cc
def test_partition_counts_reference_case(oracle, ref):
    top = np.uint64(1) << np.uint64(63)
    i = np.arange(5, dtype=np.uint64)
    keys = np.stack([i, i | top, i | top | np.uint64(1024)], axis=1).reshape(-1)
    k = Keys(u64=keys)
    c_ref = ref.partition_counts(k, 2, 1)
    assert list(c_ref) == [5, 10]
    assert np.array_equal(oracle.partition_counts(keys, 1), c_ref)


def test_partition_counts_strings(oracle, ref):
    off, chars, _, _ = datagen.strgen(20000, seed=2)
    k = Keys(offsets=off, chars=chars)
    for bits in (1, 3, 10):
        assert np.array_equal(oracle.partition_counts(oracle.hash_strings(off, chars), bits),
                              ref.partition_counts(k, 4, bits))


# ---- hashjoin.h has no reference test: pin the walk restatement on the reference itself
def _walk_vs_ref(oracle, ref, rkeys, skeys, threads=1):
    out_r, out_s, sm, _ = ref.join(rkeys, skeys, threads=threads)
    rh, rr = oracle.radix_sort(oracle.hashes(rkeys), rkeys)
    sh, sr = oracle.radix_sort(oracle.hashes(skeys), skeys)
    w_r, w_s = oracle.merge_walk_reference(rh, rr, sh, sr, rkeys, skeys)
    assert np.array_equal(w_r, out_r) and np.array_equal(w_s, out_s)
    assert int(w_r.astype(np.uint64).sum() + w_s.astype(np.uint64).sum()) == sm
    return out_r, out_s


def test_merge_walk_unique_u64(oracle, ref):
    keys = datagen.unique_u64(20000, seed=1)
    r, s = Keys(u64=datagen.permuted(keys, 1)), Keys(u64=datagen.permuted(keys, 2))
    out_r, out_s = _walk_vs_ref(oracle, ref, r, s, threads=4)
    # unique keys: the reference is exact except for pairs lost once a cursor reaches the last
    # element; the true join has n pairs.
    t_r, t_s, _ = oracle.full_join(r, s)
    assert t_r.size == 20000 and out_r.size in (19999, 20000)


def test_merge_walk_defects_documented(oracle, ref):
    # R=[a], S=[a,a,a] -> 1 of 3;  2x2 duplicates -> 3 of 4 (L-shaped walk)
    r, s = Keys(u64=np.array([7], np.uint64)), Keys(u64=np.array([7, 7, 7], np.uint64))
    assert _walk_vs_ref(oracle, ref, r, s)[0].size == 1
    assert oracle.full_join(r, s)[0].size == 3
    r, s = Keys(u64=np.array([5, 5, 9], np.uint64)), Keys(u64=np.array([5, 5, 9, 100], np.uint64))
    assert _walk_vs_ref(oracle, ref, r, s)[0].size == 4  # 3 of the 2x2 block + (9,9)
    assert oracle.full_join(r, s)[0].size == 5


def test_merge_walk_random_with_duplicates(oracle, ref):
    rng = np.random.default_rng(4)
    for n in (50, 3000):
        r = Keys(u64=rng.integers(0, n // 2, size=n, dtype=np.uint64))
        s = Keys(u64=rng.integers(0, n // 2, size=n + 17, dtype=np.uint64))
        _walk_vs_ref(oracle, ref, r, s, threads=2)


def test_merge_walk_strings(oracle, ref):
    off, chars, _, _ = datagen.strgen(15000, seed=3, shuffle_seed=1)
    off2, chars2, _, _ = datagen.strgen(15000, seed=3, shuffle_seed=2)
    r, s = Keys(offsets=off, chars=chars), Keys(offsets=off2, chars=chars2)
    out_r, _ = _walk_vs_ref(oracle, ref, r, s, threads=4)
    assert oracle.full_join(r, s)[0].size == 15000 and out_r.size >= 14999


def test_equi_join_equals_reference_where_reference_is_exact(oracle, ref):
    # unique keys + a non-matching maximal sentinel row on each side: the reference's
    # early stop cannot lose a pair, so its output must equal the true equi-join.
    keys = datagen.unique_u64(5000, seed=8) >> np.uint64(1)  # keep below the sentinels
    rk = np.concatenate([datagen.permuted(keys, 3), [np.uint64(2 ** 64 - 1)]]).astype(np.uint64)
    sk = np.concatenate([datagen.permuted(keys, 4), [np.uint64(2 ** 64 - 2)]]).astype(np.uint64)
    r, s = Keys(u64=rk), Keys(u64=sk)
    out_r, out_s, _, _ = ref.join(r, s, threads=2)
    t_r, t_s, _ = oracle.full_join(r, s)
    c_r, c_s = oracle.sort_pairs(out_r, out_s)
    assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)
