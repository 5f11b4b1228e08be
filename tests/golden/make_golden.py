"""Generates the committed golden fixtures from the UNMODIFIED reference (oracle/_ref, built by
`make ref` from /root/reference).  Run here (the reference is not on the GPU box):
    python tests/golden/make_golden.py
Fixtures are small JSON files; 64-bit values are hex strings."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from hashmergejoin_b200 import datagen  # noqa: E402
from tests.oracle_binding import Keys, load_ref  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def hx(a):
    return [format(int(x), "016x") for x in a]


def main():
    ref = load_ref()
    assert ref is not None, "run `make ref` first"
    rng = np.random.default_rng(2024)
    # 1. std::hash<std::string> vectors: fixed strings + every length 0..40 of random bytes
    items = [b"", b"a", b"abc", b"hello-world", b"aardvark-zebra", b"0123456789abcdef",
             b"the quick brown fox"]
    items += [bytes(rng.integers(0, 256, size=ln, dtype=np.uint8)) for ln in range(0, 41)]
    items += [bytes(rng.integers(97, 123, size=ln, dtype=np.uint8)) for ln in (63, 64, 65, 127, 200)]
    json.dump({"comment": "std::hash<std::string> (libstdc++ 13.3 x86-64) from oracle/_ref",
               "strings_hex": [b.hex() for b in items],
               "hashes": hx(ref.hash_bytes(b) for b in items)},
              open(os.path.join(HERE, "std_hash_strings.json"), "w"), indent=0)
    # 2. radix_non_inplace_par output (hash + row id order) for seeded inputs
    sorts = []
    for n, bits, seed in [(5, 3, 0), (257, 1, 1), (1000, 0, 2), (1000, 6, 3)]:
        keys = np.random.default_rng(seed).integers(0, 2 ** 64 if seed else 8, size=n, dtype=np.uint64)
        h, r, _ = ref.sort(Keys(u64=keys), threads=2, bits=bits)
        sorts.append({"keys": hx(keys), "bits": bits, "hash": hx(h), "rowid": [int(x) for x in r]})
    json.dump({"comment": "radix_hash::radix_non_inplace_par<uint64_t,uint64_t> from oracle/_ref",
               "cases": sorts}, open(os.path.join(HERE, "radix_sort_u64.json"), "w"), indent=0)
    # 3. strgen-shaped strings: hashes, sorted order, HashMergeJoin output
    off, chars, _, _ = datagen.strgen(600, seed=77, shuffle_seed=1)
    off2, chars2, _, _ = datagen.strgen(600, seed=77, shuffle_seed=2)
    r, s = Keys(offsets=off, chars=chars), Keys(offsets=off2, chars=chars2)
    h, rid, _ = ref.sort(r, threads=1, bits=0)
    out_r, out_s, sm, _ = ref.join(r, s, threads=1)
    json.dump({"comment": "strgen(600) x strgen(600): reference sort of R and HashMergeJoin(R,S) pairs",
               "r_offsets": [int(x) for x in off], "r_chars_hex": chars.tobytes().hex(),
               "s_offsets": [int(x) for x in off2], "s_chars_hex": chars2.tobytes().hex(),
               "r_hash_unsorted": hx(ref.hash_strings(off, chars)),
               "r_sorted_hash": hx(h), "r_sorted_rowid": [int(x) for x in rid],
               "join_r": [int(x) for x in out_r], "join_s": [int(x) for x in out_s], "join_sum": int(sm)},
              open(os.path.join(HERE, "strgen_join_600.json"), "w"), indent=0)
    # 4. u64 join with duplicates: reference walk output (defects included)
    rk = np.random.default_rng(5).integers(0, 40, size=120, dtype=np.uint64)
    sk = np.random.default_rng(6).integers(0, 40, size=150, dtype=np.uint64)
    out_r, out_s, sm, _ = ref.join(Keys(u64=rk), Keys(u64=sk), threads=1)
    json.dump({"comment": "HashMergeJoin on duplicate-heavy u64 keys: the reference walk (L-shaped, early stop)",
               "r_keys": hx(rk), "s_keys": hx(sk), "join_r": [int(x) for x in out_r],
               "join_s": [int(x) for x in out_s], "join_sum": int(sm)},
              open(os.path.join(HERE, "dup_join_u64.json"), "w"), indent=0)
    # 5. partition_only counts + optimal_partition table
    keys = np.random.default_rng(9).integers(0, 2 ** 64, size=4000, dtype=np.uint64)
    json.dump({"comment": "partition_only sizes and optimal_partition values from oracle/_ref",
               "keys": hx(keys),
               "counts": {str(b): [int(x) for x in ref.partition_counts(Keys(u64=keys), 3, b)] for b in (1, 4, 10)},
               "optimal_partition": {str(n): ref.lib.ref_optimal_partition(n) for n in
                                     [0, 1, 63, 64, 100, 4095, 4096, 10 ** 4, 10 ** 6, 2 ** 20, 10 ** 7,
                                      64 * 10 ** 6, 2 ** 26, 2 ** 28, 5 * 10 ** 8, 2 ** 29, 10 ** 9, 2 ** 30]}},
              open(os.path.join(HERE, "partition_u64.json"), "w"), indent=0)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
