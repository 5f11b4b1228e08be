"""Host-side logic of the multi-GPU path (hashmergejoin_b200/dist.py) on CPU: world_size 2 and 4
over gloo.  Each rank owns a block of both relations, splits it on the top log2(G) hash bits
(here by a numpy restatement of what K8 / hmj_partition_pairs_base produces: buckets in
destination order, stable inside a bucket, global row ids), runs the real exchange_pairs() and
then sort-merges its own hash range with the CPU oracle.  Checked:
  * concatenating the ranks' sorted hash arrays == the single-process sorted array
  * the union of the ranks' join results == the single-process join (canonical order)
  * no hash range straddles two ranks; counts matrix is consistent.
No GPU, no NCCL: the exchange code path is the same function the GPU path calls."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from hashmergejoin_b200 import datagen  # noqa: E402
from hashmergejoin_b200.dist import exchange_pairs, exchange_strings, partition_bits, plan_exchange  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def host_partition(keys, rowid_base, bits):
    """numpy restatement of hmj_partition_pairs_base: stable grouping on hash >> (64-bits)"""
    n = keys.size
    dest = (keys >> np.uint64(64 - bits)).astype(np.int64) if bits else np.zeros(n, np.int64)
    order = np.argsort(dest, kind="stable")
    counts = np.bincount(dest, minlength=1 << bits)
    offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    rowid = (np.arange(n, dtype=np.uint32) + np.uint32(rowid_base))
    return keys[order], rowid[order], offs


def _worker(rank, world, port, n_total, dup, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import sys
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from tests.oracle_binding import Keys, load_oracle
        oracle = load_oracle()
        bits = partition_bits(world)
        rk_all, sk_all = make_relations(n_total, dup)
        n = n_total // world
        side = {}
        for name, allk in (("r", rk_all), ("s", sk_all)):
            mine = allk[rank * n:(rank + 1) * n]
            ph, pr, offs = host_partition(mine, rank * n, bits)
            t = lambda a, dt: torch.from_numpy(a.view(dt).copy())
            rh, rr, mat = exchange_pairs(t(ph, np.int64), t(pr, np.int32), torch.from_numpy(offs), rank, world)
            send, recv = plan_exchange(mat, rank)
            assert sum(recv) == rh.numel() and sum(send) == n
            h = rh.numpy().view(np.uint64)
            r = rr.numpy().view(np.uint32)
            if bits:
                assert np.all((h >> np.uint64(64 - bits)) == rank), "received a row of another range"
            order = np.argsort(h, kind="stable")  # local sort of my hash range (oracle: std::sort order)
            side[name] = (h[order], r[order])
        (rh, rr), (sh, sr) = side["r"], side["s"]
        # local merge join of my range with the oracle's true equi-join on (hash,row id) streams
        out_r, out_s, _ = oracle.equi_join(rh, rr, sh, sr, Keys(u64=rk_all), Keys(u64=sk_all))
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), rh=rh, rr=rr, sh=sh, sr=sr, out_r=out_r, out_s=out_s)
    finally:
        dist.destroy_process_group()


def make_relations(n_total, dup):
    keys = datagen.unique_u64(n_total // dup, seed=7)
    rk = datagen.permuted(np.tile(keys, dup), 1)
    sk = datagen.permuted(np.tile(keys, dup), 2)
    return rk, sk


@pytest.mark.parametrize("world,dup", [(2, 1), (2, 3), (4, 1)])
def test_dist_join_host_logic_gloo(tmp_path, oracle, world, dup):
    from tests.oracle_binding import Keys
    n_total = 6144 * dup
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_total, dup, str(tmp_path)), nprocs=world, join=True)
    rk, sk = make_relations(n_total, dup)
    parts = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    # (1) rank-order concatenation of the sorted hash arrays == global sorted array
    assert np.array_equal(np.concatenate([p["rh"] for p in parts]), np.sort(rk))
    assert np.array_equal(np.concatenate([p["sh"] for p in parts]), np.sort(sk))
    # row ids are global and complete
    assert np.array_equal(np.sort(np.concatenate([p["rr"] for p in parts])), np.arange(n_total, dtype=np.uint32))
    # (2) union of per-rank results == single-process join, canonical order
    t_r, t_s, _ = oracle.full_join(Keys(u64=rk), Keys(u64=sk))
    g_r = np.concatenate([p["out_r"] for p in parts])
    g_s = np.concatenate([p["out_s"] for p in parts])
    c_r, c_s = oracle.sort_pairs(g_r, g_s)
    assert c_r.size == n_total * dup
    assert np.array_equal(c_r, t_r) and np.array_equal(c_s, t_s)


def make_string_relations(n_keys):
    """C5's shape in small: build side holds every key twice, probe side draws keys Zipf(1.0)"""
    off, chars, _, _ = datagen.strgen(n_keys, seed=31, shuffle_seed=3)
    rng = np.random.default_rng(17)
    r = datagen.take_rows(off, chars, rng.permutation(np.repeat(np.arange(n_keys), 2)))
    s = datagen.take_rows(off, chars, datagen.zipf_indices(2 * n_keys, n_keys, 1.0, seed=5))
    return r, s


def _string_worker(rank, world, port, n_keys, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import sys
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from tests.oracle_binding import Keys, load_oracle
        oracle = load_oracle()
        bits = partition_bits(world)
        rel = make_string_relations(n_keys)
        n = 2 * n_keys // world
        got = {}
        for name, (off, chars) in zip("rs", rel):
            m_off, m_chars = datagen.take_rows(off, chars, np.arange(rank * n, (rank + 1) * n))
            # what the device does before the exchange: hash, group rows by destination (stable),
            # put the keys in that order
            h = oracle.hash_strings(m_off, m_chars)
            _, rows, offs = host_partition(h, 0, bits)
            k_off, k_chars = datagen.take_rows(m_off, m_chars, rows.astype(np.int64))
            t = torch.from_numpy
            rid, r_off, r_chars, sent = exchange_strings(
                t((rows + np.uint32(rank * n)).view(np.int32).copy()), t(offs), t(k_off.view(np.int64).copy()),
                t(k_chars.copy()), rank, world)
            r_off, r_chars = r_off.numpy().view(np.uint64), r_chars.numpy()
            assert r_off[-1] == r_chars.size and sent > 0
            hh = oracle.hash_strings(r_off, r_chars)
            assert np.all((hh >> np.uint64(64 - bits)) == rank), "received a key of another hash range"
            got[name] = (rid.numpy().view(np.uint32), r_off, r_chars)
        (r_id, r_off, r_chars), (s_id, s_off, s_chars) = got["r"], got["s"]
        # local join of my range on arrival positions, then positions -> global rows
        t_r, t_s, _ = oracle.full_join(Keys(offsets=r_off, chars=r_chars), Keys(offsets=s_off, chars=s_chars))
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), out_r=r_id[t_r], out_s=s_id[t_s])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_dist_string_join_host_logic_gloo(tmp_path, oracle, world):
    from tests.oracle_binding import Keys
    n_keys = 3000
    mp.spawn(_string_worker, args=(world, _free_port(), n_keys, str(tmp_path)), nprocs=world, join=True)
    (roff, rchars), (soff, schars) = make_string_relations(n_keys)
    parts = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    t_r, t_s, _ = oracle.full_join(Keys(offsets=roff, chars=rchars), Keys(offsets=soff, chars=schars))
    c_r, c_s = oracle.sort_pairs(np.concatenate([p["out_r"] for p in parts]),
                                 np.concatenate([p["out_s"] for p in parts]))
    e_r, e_s = oracle.sort_pairs(t_r, t_s)
    assert c_r.size == 4 * n_keys
    assert np.array_equal(c_r, e_r) and np.array_equal(c_s, e_s)


def test_plan_exchange_and_bits():
    mat = [[1, 2, 3, 4], [5, 6, 7, 8], [9, 10, 11, 12], [13, 14, 15, 16]]
    send, recv = plan_exchange(mat, 2)
    assert send == [9, 10, 11, 12] and recv == [3, 7, 11, 15]
    assert [partition_bits(g) for g in (1, 2, 4, 8)] == [0, 1, 2, 3]
    with pytest.raises(AssertionError):
        partition_bits(3)
