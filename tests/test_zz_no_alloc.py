"""hmj_partition_pairs_noalloc (include/hmj.h): the stable partition with its scratch in a fixed
device array inside the library, for callers that may not call cudaMalloc.  No counterpart in
the reference, so the checker is the existing entry
point hmj_partition_pairs_base on the same inputs: bit-exact (integer work), on calls in a row
that share the array and leave different contents behind.

In a file of its own, collected last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

MAX_ROWS = 1 << 24  # HMJ_NOALLOC_MAX_ROWS


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _keys(n, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    return torch.randint(-2 ** 63, 2 ** 63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)


def _outputs(n, bits):
    return (torch.full((n,), -1, dtype=torch.int64, device="cuda"),
            torch.full((n,), -1, dtype=torch.int32, device="cuda"),
            torch.full(((1 << bits) + 1,), -1, dtype=torch.int64, device="cuda"))


def _call(ctx, name, keys, rowid, base, bits, out):
    fn = getattr(ctx.lib, name)
    rc = fn(ctx.h, _p(keys), _p(rowid), base, keys.numel(), bits, _p(out[0]), _p(out[1]), _p(out[2]))
    return rc


def _err(ctx):
    return ctx.lib.hmj_last_error(ctx.h).decode()


def _same(a, b):
    return all(torch.equal(x, y) for x, y in zip(a, b))


def test_noalloc_partition_matches_the_allocating_entry_point_on_calls_in_a_row(ctx):
    # (rows, bits, explicit row ids, row id base): sizes go up and down so that every call finds
    # the previous call's histogram, counter and descriptors in the shared array
    cases = [(1_000_003, 8, False, 0), (70_001, 3, True, 0), (3_000_017, 5, False, 12345),
             (3_000_017, 5, False, 12345), (4096, 1, False, 7), (1, 8, False, 0), (0, 4, False, 0),
             (2_500_000, 8, True, 0)]
    # set-up outside the measured calls: both entry points once (workspace of the old one, symbol
    # address of the new one)
    k0 = _keys(3_000_017, 99)
    assert _call(ctx, "hmj_partition_pairs_base", k0, None, 0, 8, _outputs(k0.numel(), 8)) == 0
    assert _call(ctx, "hmj_partition_pairs_noalloc", k0[:0], None, 0, 8, _outputs(0, 8)) == 0
    ctx.synchronize()
    for i, (n, bits, with_ids, base) in enumerate(cases):
        keys = _keys(n, 1000 + i)
        if n > 16:
            keys[: n // 3] |= -2 ** 63          # uneven partitions
            keys[5] = keys[11]                  # equal hashes keep their input order
        rowid = torch.randperm(n, device="cuda").to(torch.int32) if with_ids else None
        want, got = _outputs(n, bits), _outputs(n, bits)
        if n:
            assert _call(ctx, "hmj_partition_pairs_base", keys, rowid, base, bits, want) == 0, _err(ctx)
        else:
            want[2].zero_()  # (the old entry point wants non-null columns even for no rows)
        ctx.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        launches_before = ctx.launch_count()
        assert _call(ctx, "hmj_partition_pairs_noalloc", keys, rowid, base, bits, got) == 0, _err(ctx)
        ctx.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the GPU's free memory"
        assert ctx.launch_count() > launches_before
        assert _same(want, got), f"case {i}: n={n} bits={bits}"
        # and the result is what the entry point promises, independent of the other path
        offs = got[2].cpu().numpy()
        assert offs[0] == 0 and offs[-1] == n and np.all(np.diff(offs) >= 0)
        if n:
            part = (got[0].cpu().numpy().view(np.uint64) >> np.uint64(64 - bits)).astype(np.int64)
            assert np.all(np.diff(part) >= 0)
            assert np.array_equal(np.bincount(part, minlength=1 << bits), np.diff(offs))


def test_noalloc_partition_at_the_limit_and_beyond(ctx):
    keys = _keys(MAX_ROWS, 5)
    want, got = _outputs(MAX_ROWS, 8), _outputs(MAX_ROWS, 8)
    assert _call(ctx, "hmj_partition_pairs_base", keys, None, 0, 8, want) == 0
    for _ in range(2):  # two calls in a row on the full array
        got[0].fill_(-1)
        got[1].fill_(-1)
        assert _call(ctx, "hmj_partition_pairs_noalloc", keys, None, 0, 8, got) == 0, _err(ctx)
        ctx.synchronize()
        assert _same(want, got)
    del want, got
    big = torch.empty(MAX_ROWS + 1, dtype=torch.int64, device="cuda")
    out = _outputs(MAX_ROWS + 1, 8)
    rc = _call(ctx, "hmj_partition_pairs_noalloc", big, None, 0, 8, out)
    assert rc != 0
    msg = _err(ctx)
    assert "HMJ_NOALLOC_MAX_ROWS" in msg and "16777216" in msg
    assert int(out[2][0]) == -1, "a rejected call must not touch the outputs"
    # the context and the array stay usable after the rejection
    small = _keys(100_000, 6)
    want, got = _outputs(100_000, 4), _outputs(100_000, 4)
    assert _call(ctx, "hmj_partition_pairs_base", small, None, 0, 4, want) == 0
    assert _call(ctx, "hmj_partition_pairs_noalloc", small, None, 0, 4, got) == 0
    ctx.synchronize()
    assert _same(want, got)
    # the host-side mirror binds the same symbol
    got2 = ctx.partition_pairs_noalloc(small, None, 4, _outputs(100_000, 4))
    ctx.synchronize()
    assert _same(want, got2)
