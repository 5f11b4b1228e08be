"""The CPU oracle against the committed golden fixtures (tests/golden/*.json, generated from the
unmodified reference by tests/golden/make_golden.py).  CPU only; needs no reference."""
import numpy as np

from tests import golden_util as G
from tests.oracle_binding import Keys


def test_std_hash_golden(oracle):
    g = G.load("std_hash_strings.json")
    for s, h in zip(g["strings_hex"], g["hashes"]):
        assert oracle.hash_bytes(bytes.fromhex(s)) == int(h, 16)


def test_radix_sort_golden(oracle):
    for case in G.load("radix_sort_u64.json")["cases"]:
        keys = G.u64(case["keys"])
        h, r = oracle.radix_sort(keys, Keys(u64=keys), bits=case["bits"])
        assert np.array_equal(h, G.u64(case["hash"]))
        assert np.array_equal(r, np.array(case["rowid"], np.uint32))


def test_strgen_join_golden(oracle):
    g = G.load("strgen_join_600.json")
    r = Keys(offsets=np.array(g["r_offsets"], np.uint64), chars=G.u8(g["r_chars_hex"]))
    s = Keys(offsets=np.array(g["s_offsets"], np.uint64), chars=G.u8(g["s_chars_hex"]))
    rh_u = oracle.hashes(r)
    assert np.array_equal(rh_u, G.u64(g["r_hash_unsorted"]))
    rh, rr = oracle.radix_sort(rh_u, r)
    assert np.array_equal(rh, G.u64(g["r_sorted_hash"]))
    assert np.array_equal(rr, np.array(g["r_sorted_rowid"], np.uint32))
    sh, sr = oracle.radix_sort(oracle.hashes(s), s)
    w_r, w_s = oracle.merge_walk_reference(rh, rr, sh, sr, r, s)
    assert np.array_equal(w_r, np.array(g["join_r"], np.uint32))
    assert np.array_equal(w_s, np.array(g["join_s"], np.uint32))
    # unique keys: the true equi-join contains the reference's pairs and at most one more
    t_r, t_s, _ = oracle.full_join(r, s)
    ref_pairs = set(zip(g["join_r"], g["join_s"]))
    true_pairs = set(zip(t_r.tolist(), t_s.tolist()))
    assert ref_pairs <= true_pairs and len(true_pairs) == 600 and len(true_pairs - ref_pairs) <= 1


def test_dup_join_golden(oracle):
    g = G.load("dup_join_u64.json")
    r, s = Keys(u64=G.u64(g["r_keys"])), Keys(u64=G.u64(g["s_keys"]))
    rh, rr = oracle.radix_sort(r.u64, r)
    sh, sr = oracle.radix_sort(s.u64, s)
    w_r, w_s = oracle.merge_walk_reference(rh, rr, sh, sr, r, s)
    assert np.array_equal(w_r, np.array(g["join_r"], np.uint32))
    assert np.array_equal(w_s, np.array(g["join_s"], np.uint32))
    # the true equi-join is the brute-force cross product
    t_r, t_s, sm = oracle.full_join(r, s)
    brute = sorted((i, j) for i, a in enumerate(r.u64) for j, b in enumerate(s.u64) if a == b)
    assert list(zip(t_r.tolist(), t_s.tolist())) == brute
    assert sm == sum(i + j for i, j in brute)
    assert set(zip(g["join_r"], g["join_s"])) <= set(brute)


def test_partition_golden(oracle):
    g = G.load("partition_u64.json")
    keys = G.u64(g["keys"])
    for bits, counts in g["counts"].items():
        assert oracle.partition_counts(keys, int(bits)).tolist() == counts
    for n, k in g["optimal_partition"].items():
        assert oracle.optimal_partition(int(n)) == k


def test_hash_join_raw_semantics(oracle):
    # hashjoin_bench.cc:47-53: last S row wins for duplicate keys, misses add r.value only
    r = Keys(u64=np.array([1, 2, 3, 9], np.uint64))
    s = Keys(u64=np.array([2, 3, 2, 7], np.uint64))
    hits, sm = oracle.hash_join_raw(r, s)
    assert hits == 2
    assert sm == (0 + 1 + 2 + 3) + 2 + 1  # r rows + s_map[2]=row 2 + s_map[3]=row 1
