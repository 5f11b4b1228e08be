"""The bucket plan drawn from a sample (bucket_join.cu: hmj_plan_from_sample), checked on the host:
for a set of key distributions the plan's capacities are compared with the occupancy the scatter
levels would really see (numpy on the full key set).  No GPU needed: the function is host code,
reached through hmj_debug_plan_from_sample."""
import ctypes as C
import math

import numpy as np
import pytest

from hashmergejoin_b200 import _lib

PER_SIDE = 32768
CAP_SEG = 4608


def sample(keys):
    n = keys.size
    idx = (np.arange(PER_SIDE, dtype=np.uint64) * np.uint64(n) // np.uint64(PER_SIDE)).astype(np.int64)
    return keys[idx]


def plan(rk, sk):
    lib = _lib.load()
    h = np.ascontiguousarray(np.concatenate([sample(rk), sample(sk)]), dtype=np.uint64)
    out = np.zeros(10, np.int64)
    rc = lib.hmj_debug_plan_from_sample(C.c_void_p(h.ctypes.data), PER_SIDE, rk.size, sk.size,
                                        C.c_void_p(out.ctypes.data))
    assert rc == 0
    names = ["levels", "bits0", "bits1", "qb", "skip", "cap_seg", "check_prefix", "prefix", "density6", "skewed"]
    return dict(zip(names, (int(x) for x in out)))


def occupancy_ok(p, keys):
    """would the scatter levels of plan p hold `keys`?  (hmj_uniform_scatter's capacities)"""
    n = keys.size
    dens = p["density6"] / 1e6
    if p["skip"]:
        if not np.all((keys >> np.uint64(64 - p["skip"])) == np.uint64(p["prefix"] & (2 ** 64 - 1))):
            return False
    sub = (keys << np.uint64(p["skip"])) if p["skip"] else keys
    if p["levels"] == 2:
        avg1 = n * dens / (1 << p["bits0"])
        cap1 = (int(avg1 + 8.0 * math.sqrt(avg1) + 64.0) + 255) & ~255
        c1 = np.bincount((sub >> np.uint64(64 - p["bits0"])).astype(np.int64), minlength=1 << p["bits0"])
        if c1.max() > cap1:
            return False
    c2 = np.bincount((sub >> np.uint64(64 - p["qb"])).astype(np.int64), minlength=1 << p["qb"])
    return c2.max() <= p["cap_seg"]


def splitmix(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def test_uniform_hashes_keep_the_plain_plan():
    rng = np.random.default_rng(1)
    for n in (1 << 22, 5_000_000, 1 << 24):
        with np.errstate(over="ignore"):
            rk = splitmix(rng.permutation(n).astype(np.uint64))
            sk = splitmix(rng.permutation(n).astype(np.uint64))
        p = plan(rk, sk)
        qb = max(0, math.ceil(math.log2(n / 4096)))
        assert p["levels"] == 2 and p["qb"] == qb and p["skip"] == 0 and p["density6"] == 1_000_000
        assert p["skewed"] == 0 and p["cap_seg"] == CAP_SEG
        assert occupancy_ok(p, rk) and occupancy_ok(p, sk)


@pytest.mark.parametrize("n", [1 << 22, 1 << 23])
def test_dense_power_of_two_range_is_a_prefix_only(n):
    rng = np.random.default_rng(2)
    rk, sk = rng.permutation(n).astype(np.uint64), rng.permutation(n).astype(np.uint64)
    p = plan(rk, sk)
    assert p["skip"] == 64 - int(math.log2(n)) and p["check_prefix"] == 1 and p["prefix"] == 0
    assert p["density6"] == 1_000_000 and p["qb"] == int(math.log2(n)) - 12 and p["skewed"] == 0
    assert occupancy_ok(p, rk) and occupancy_ok(p, sk)


@pytest.mark.parametrize("n,base", [(3 << 20, 0), (5_000_000, 0), (6_291_456, 0), ((1 << 22) + 1000, 0),
                                    (7_654_321, 0), (3 << 21, 1 << 40), (5_000_000, (1 << 33) + 12345),
                                    (4_500_000, (1 << 23) - 2_000_000)])
def test_dense_range_that_is_not_a_power_of_two_gets_a_density(n, base):
    rng = np.random.default_rng(3)
    rk = rng.permutation(n).astype(np.uint64) + np.uint64(base)
    sk = rng.permutation(n).astype(np.uint64) + np.uint64(base)
    p = plan(rk, sk)
    assert p["levels"] == 2 and p["skewed"] == 0, p
    lo, hi = int(rk.min()), int(rk.max())
    k = (lo ^ hi).bit_length()
    assert p["skip"] == 64 - k and p["check_prefix"] == 1
    true_density = 2 ** k / n
    assert abs(p["density6"] / 1e6 - true_density) < 2e-3 * true_density
    assert p["qb"] == k - 12, "one bucket per 4096 key values"
    assert occupancy_ok(p, rk) and occupancy_ok(p, sk)


def test_sides_of_different_size_and_range():
    rng = np.random.default_rng(4)
    n = 6_000_000
    rk = rng.permutation(n).astype(np.uint64)
    # S: a third of the rows, spread over the whole range / packed into its first third
    sk_spread = rng.choice(n, size=n // 3, replace=False).astype(np.uint64)
    sk_packed = rng.permutation(n // 3).astype(np.uint64)
    for sk in (sk_spread, sk_packed):
        for a, b in ((rk, sk), (sk, rk)):
            p = plan(a, b)
            if p["levels"] and not p["skewed"]:
                assert occupancy_ok(p, a) and occupancy_ok(p, b), p


def test_too_sparse_or_duplicated_dense_keys_are_left_to_the_old_checks():
    rng = np.random.default_rng(5)
    # a range 9x smaller than the space under its prefix: no density plan (the skew check speaks)
    n = 1 << 22
    rk = rng.permutation(n).astype(np.uint64) + np.uint64((1 << 25) - (n // 2))
    p = plan(rk, rk[::-1].copy())
    assert p["density6"] == 1_000_000
    assert p["skewed"] != 0 or not occupancy_ok(p, rk)
    # every dense key three times (3 rows per key value): 4096 values would not fit a bucket
    n = 5_000_000
    rk3 = np.repeat(rng.permutation(n // 3).astype(np.uint64), 3)
    rng.shuffle(rk3)
    p = plan(rk3, rk3[::-1].copy())
    assert p["density6"] == 1_000_000


def test_zipf_heads_are_still_called_skewed():
    rng = np.random.default_rng(6)
    n = 1 << 22
    with np.errstate(over="ignore"):
        uniq = splitmix(np.arange(n, dtype=np.uint64))
    ranks = np.minimum(rng.zipf(1.2, size=n) - 1, n - 1)
    sk = uniq[ranks]
    p = plan(uniq, sk)
    assert p["skewed"] & 2
