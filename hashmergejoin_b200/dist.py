"""Multi-GPU hash-merge join: one process per GPU, hash-range partitioning, ONE exchange.

Replaces the reference's thread-level partitioning (partition_only, /root/reference
partitioned_hash.h:82-124, and the per-thread slices of radix_hash.h:375-388) with the scheme of
SURVEY.md section 8(e): every rank splits its rows of R and S on the top log2(G) hash bits
(K8, hmj_partition_pairs_base -- destination g = hash >> (64-b), exactly the reference's
partition function), the G x G counts are all-gathered, one all-to-all moves bucket g to rank g,
and each rank then sort-merges its own hash range locally (K2-K6).  Ranges are ordered by rank, so
concatenating the per-rank sorted arrays gives the single-GPU sorted array, and equal hashes
never straddle ranks.  torch.distributed is plumbing only: NCCL all_to_all_single on the stream
the kernels run on (gloo on CPU tensors in the world_size-2 host-logic tests).
"""
import math
import os

import torch
import torch.distributed as dist

from ._lib import HmjError
from .api import SortedRelation


def partition_bits(world):
    assert world >= 1 and world & (world - 1) == 0, \
        "hash-range partitioning needs a power-of-two GPU count"
    return int(math.log2(world))


def plan_exchange(counts_matrix, rank):
    """counts_matrix[src][dst] = rows going src->dst.  Returns (send_splits, recv_splits)."""
    send = [int(x) for x in counts_matrix[rank]]
    recv = [int(counts_matrix[src][rank]) for src in range(len(counts_matrix))]
    return send, recv


def exchange_pairs(part_hash, part_rowid, part_offsets, rank, world, group=None):
    """The one exchange step.  part_hash/part_rowid are this rank's pairs grouped by destination
    (bucket g = rows whose hash >> (64-b) == g, at [part_offsets[g], part_offsets[g+1])).
    Returns (hash, rowid) of everything destined to this rank, source-rank-major, plus the
    G x G count matrix.  Works on whatever device the tensors live on."""
    counts = (part_offsets[1:world + 1] - part_offsets[:world]).to(torch.int64).contiguous()
    allc = torch.empty(world * world, dtype=torch.int64, device=counts.device)
    dist.all_gather_into_tensor(allc, counts, group=group)
    mat = allc.view(world, world).cpu().tolist()
    send, recv = plan_exchange(mat, rank)
    assert sum(send) == part_hash.numel(), "partition offsets do not cover the local rows"
    nrecv = sum(recv)
    rh = torch.empty(nrecv, dtype=part_hash.dtype, device=part_hash.device)
    rr = torch.empty(nrecv, dtype=part_rowid.dtype, device=part_rowid.device)
    dist.all_to_all_single(rh, part_hash, recv, send, group=group)
    dist.all_to_all_single(rr, part_rowid, recv, send, group=group)
    return rh, rr, mat


def exchange_strings(part_rowid, part_offsets, key_offsets, key_chars, rank, world, group=None, mark=None,
                     key_lens=None):
    """The exchange step for string keys: the keys travel with their rows.
    part_rowid: this rank's global row ids grouped by destination (bucket g at
    [part_offsets[g], part_offsets[g+1])); key_offsets/key_chars: the keys in that same order.
    Returns (rowid, offsets, chars) of everything destined to this rank, source-rank-major, and
    the bytes this rank sent to other ranks.  The hashes stay behind: the receiver hashes the
    keys again (8 bytes per row less on the wire).  Works on whatever device the tensors are on."""
    dev = part_rowid.device
    po = part_offsets[:world + 1].to(torch.int64)
    rows = (po[1:] - po[:-1]).contiguous()
    ko = key_offsets.to(torch.int64)
    nbytes = (ko[po[1:]] - ko[po[:-1]]).contiguous()
    mine = torch.cat([rows, nbytes])
    allc = torch.empty(world * 2 * world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(allc, mine, group=group)
    allc = allc.view(world, 2, world).cpu()
    row_mat, byte_mat = allc[:, 0, :].tolist(), allc[:, 1, :].tolist()
    send_rows, recv_rows = plan_exchange(row_mat, rank)
    send_bytes, recv_bytes = plan_exchange(byte_mat, rank)
    assert sum(send_rows) == part_rowid.numel() and sum(send_bytes) == key_chars.numel(), \
        "partition offsets do not cover the local rows"
    lens = key_lens if key_lens is not None else (ko[1:] - ko[:-1]).to(torch.int32).contiguous()
    r_rowid = torch.empty(sum(recv_rows), dtype=part_rowid.dtype, device=dev)
    r_lens = torch.empty(sum(recv_rows), dtype=torch.int32, device=dev)
    r_chars = torch.empty(sum(recv_bytes), dtype=key_chars.dtype, device=dev)
    if mark:
        mark("  counts + planning")
    dist.all_to_all_single(r_rowid, part_rowid.contiguous(), recv_rows, send_rows, group=group)
    dist.all_to_all_single(r_lens, lens, recv_rows, send_rows, group=group)
    if mark:
        mark("  all_to_all row ids + lengths")
    dist.all_to_all_single(r_chars, key_chars.contiguous(), recv_bytes, send_bytes, group=group)
    if mark:
        mark("  all_to_all key bytes")
    r_off = torch.zeros(r_lens.numel() + 1, dtype=torch.int64, device=dev)
    if r_lens.numel():
        torch.cumsum(r_lens, 0, dtype=torch.int64, out=r_off[1:])
    sent = 8 * (sum(send_rows) - send_rows[rank]) + sum(send_bytes) - send_bytes[rank]
    return r_rowid, r_off, r_chars, sent


class DistJoin:
    """One per rank.  join_u64 / join_strings are collective: every rank of the job calls them in
    the same order.  With ``rows_total_hint`` and the NCCL backend the integer-key join runs the
    fused path (hmj_dist_join_u64: level 1 of the bucket scatter writes into the owners' memory
    over NVLink, relations pipelined, no host barrier); otherwise, and whenever a region overflows
    (skewed keys), the NCCL path: stable partition pass + all_to_all_single + local join."""

    def __init__(self, ctx, rank, world, rows_total_hint=0, string_rows_hint=0, key_bytes_hint=24):
        self.ctx, self.rank, self.world = ctx, rank, world
        self.bits = partition_bits(world)
        self.exchange_bytes = 0      # bytes this rank sent to OTHER ranks in the last join
        self.exchange_measured = False  # read back from the device cursors (fused path), not a formula
        self.local_rows = (0, 0)     # (|R|, |S|) of this rank's hash range in the last join
        self.fused = False
        self.fused_joins = 0
        self.fallback_joins = 0
        self.trace = os.environ.get("HMJ_DIST_TRACE", "0") != "0"
        self.trace_log = []
        # string keys: receive buffers in every rank's memory that the peers' gather kernels write
        # into (hmj_xchg_*); sized from the hint with room for the head key's owner of a Zipf side
        self.xchg = False
        self.xchg_joins = 0
        self.rows_cap = self.bytes_cap = 0
        if world > 1 and string_rows_hint > 0 and dist.get_backend() == "nccl" \
                and os.environ.get("HMJ_DIST_FUSED", "1") != "0":
            self.rows_cap = int(1.6 * string_rows_hint / world) + 4096
            self.bytes_cap = self.rows_cap * key_bytes_hint
            try:
                mine = torch.from_numpy(ctx.xchg_setup(rank, world, self.rows_cap, self.bytes_cap)).cuda()
                allh = torch.empty(world * mine.numel(), dtype=torch.uint8, device=mine.device)
                dist.all_gather_into_tensor(allh, mine)
                ctx.xchg_connect(allh.cpu().numpy())
                self.xchg = True
            except Exception as e:
                self.xchg_error = str(e)
            flag = torch.tensor([1 if self.xchg else 0], device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            self.xchg = bool(flag.item())
            self._token = torch.zeros(1, dtype=torch.int32, device="cuda")
        if world > 1 and rows_total_hint > 0 and dist.get_backend() == "nccl" \
                and os.environ.get("HMJ_DIST_FUSED", "1") != "0":
            try:
                mine = torch.from_numpy(ctx.dist_setup(rank, world, rows_total_hint)).cuda()
                allh = torch.empty(world * mine.numel(), dtype=torch.uint8, device=mine.device)
                dist.all_gather_into_tensor(allh, mine)
                ctx.dist_connect(allh.cpu().numpy())
                self.fused = True
            except Exception as e:  # size outside the plan's range: the NCCL path serves it
                self.fused_error = str(e)
            flag = torch.tensor([1 if self.fused else 0], device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            self.fused = bool(flag.item())

    def close(self):
        """Collective: no rank may free its regions while another still writes into them."""
        if self.fused or self.xchg:
            torch.cuda.synchronize()
            dist.barrier()
        if self.fused:
            self.ctx.dist_teardown()
            self.fused = False
        if self.xchg:
            self.ctx.xchg_teardown()
            self.xchg = False

    def _exchange_side(self, keys, rowid_base):
        """partition local rows by destination, exchange -> (hash, rowid) tensors of my range"""
        ctx, G = self.ctx, self.world
        if G == 1:
            n = keys.numel()
            rid = torch.arange(rowid_base, rowid_base + n, dtype=torch.int64, device=keys.device).to(torch.int32)
            return keys, rid
        # bits >= 1 here; K8 groups the pairs by destination and generates global row ids
        ph, pr, offs = ctx.partition_pairs(keys, None, self.bits, rowid_base=rowid_base)
        rh, rr, mat = exchange_pairs(ph, pr, offs, self.rank, G)
        self.exchange_bytes += 12 * (sum(mat[self.rank]) - mat[self.rank][self.rank])
        return rh, rr

    def join_u64(self, rkeys_local, skeys_local, want_pairs=False, measure_exchange=False):
        """Returns the number of pairs this rank produced (its hash range).  Global row ids are
        rank-major blocks: rank r owns rows [r*n, (r+1)*n) of a relation, so every rank must pass
        blocks of the same size (checked across ranks on the fused path)."""
        ctx = self.ctx
        nr, ns = rkeys_local.numel(), skeys_local.numel()
        self.exchange_bytes = 0
        self.exchange_measured = False
        if self.fused and not want_pairs:
            # a rank that cannot even start (bad argument) must not leave the others waiting for
            # its flags until they time out: it reports not-ok and everybody takes the NCCL path
            m, ok, err = 0, 0, None
            try:
                m, ok = ctx.dist_join_u64(rkeys_local, self.rank * nr, skeys_local, self.rank * ns)
            except Exception as e:
                err = e
            # one small all-reduce: everybody ok?  same block sizes everywhere?
            t = torch.tensor([ok, nr, -nr, ns, -ns], dtype=torch.int64, device=rkeys_local.device)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            t = t.tolist()
            if t[1] != -t[2] or t[3] != -t[4]:
                raise ValueError("DistJoin.join_u64: ranks passed blocks of different sizes "
                                 "(global row ids are rank * block size)")
            if t[0] == 1:
                self.fused_joins += 1
                if measure_exchange:
                    sent, got_r, got_s = ctx.dist_last_exchange()
                    self.exchange_bytes, self.exchange_measured = 12 * sent, True
                    self.local_rows = (got_r, got_s)
                else:
                    # uniform hashes: (G-1)/G of the local rows leave the rank, 12 B per pair
                    self.exchange_bytes = 12 * (nr + ns) * (self.world - 1) // self.world
                    self.local_rows = (nr, ns)
                return m
            self.fallback_joins += 1
            if err is not None and not isinstance(err, HmjError):
                raise err
        # global row ids: rank-major blocks (rank r owns rows [r*n, (r+1)*n) of each relation)
        rh, rr = self._exchange_side(rkeys_local, self.rank * nr)
        sh, sr = self._exchange_side(skeys_local, self.rank * ns)
        self.local_rows = (rh.numel(), sh.numel())
        self.exchange_measured = True
        if want_pairs:
            # staged API (tests: sorted ranges of the ranks concatenate to the global sorted array)
            rh, rr = ctx.sort_pairs(rh, rr, out_hash=rh, out_rowid=rr)
            sh, sr = ctx.sort_pairs(sh, sr, out_hash=sh, out_rowid=sr)
            out_r, out_s = ctx.merge_join(SortedRelation(rh, rr), SortedRelation(sh, sr))
            return out_r.numel(), (out_r, out_s, rh, sh)
        # the rank's hash range shares its top log2(G) bits: bucket join below them
        return ctx.join_pairs(rh, rr, sh, sr, prefix_bits=self.bits)

    def _rank_barrier(self):
        """stream-ordered: work enqueued after it runs once every rank's earlier work is done"""
        dist.all_reduce(self._token)

    def _exchange_side_strings(self, offsets, chars, rowid_base, side=0):
        """hash local keys, group rows by destination, move row ids + lengths + key bytes to the
        owners -> (global row ids, offsets, chars) of my hash range.  Returns None for the columns
        when the fused route was taken (they are read from the receive buffers after the barrier)."""
        ctx, G = self.ctx, self.world
        n = offsets.numel() - 1
        if G == 1:
            rid = torch.arange(rowid_base, rowid_base + n, dtype=torch.int64, device=offsets.device).to(torch.int32)
            return rid, offsets, chars
        mark = self._mark
        mark("start")
        h = ctx.hash_strings(offsets, chars)
        mark("hash")
        _, rows, offs = ctx.partition_pairs(h, None, self.bits, rowid_base=0)   # local row ids by destination
        del h
        mark("partition")
        start, lens = ctx.string_row_spans(offsets, rows)
        k_off = torch.zeros(n + 1, dtype=torch.int64, device=offsets.device)
        if n:
            torch.cumsum(lens, 0, dtype=torch.int64, out=k_off[1:])
        mark("row spans + send offsets")
        # G x G matrices of rows and key bytes (source-major), the same on every rank
        po = offs[:G + 1].to(torch.int64)
        mine = torch.cat([po[1:] - po[:-1], k_off[po[1:]] - k_off[po[:-1]]]).contiguous()
        allc = torch.empty(G * 2 * G, dtype=torch.int64, device=offsets.device)
        dist.all_gather_into_tensor(allc, mine)
        allc = allc.view(G, 2, G).cpu()
        row_mat, byte_mat = allc[:, 0, :], allc[:, 1, :]
        recv_rows, recv_bytes = row_mat.sum(0), byte_mat.sum(0)   # per owner
        mark("counts")
        rid_send = rows + rowid_base
        sent = int(8 * (row_mat[self.rank].sum() - row_mat[self.rank][self.rank]) +
                   byte_mat[self.rank].sum() - byte_mat[self.rank][self.rank])
        self.exchange_bytes += sent
        if self.xchg and int(recv_rows.max()) <= self.rows_cap and int(recv_bytes.max()) <= self.bytes_cap:
            part_first = torch.zeros(G + 1, dtype=torch.int64)
            part_first[1:] = torch.cumsum(row_mat[self.rank], 0)
            row_base = row_mat[:self.rank].sum(0)      # blocks of the ranks before me, per owner
            byte_base = byte_mat[:self.rank].sum(0)
            ctx.xchg_send(side, chars, start, k_off, rid_send, lens, part_first.numpy(), row_base.numpy(),
                          byte_base.numpy())
            mark("gather + send over NVLink")
            return ("fused", int(recv_rows[self.rank]), int(recv_bytes[self.rank]))
        # NCCL route: keys into send order locally, then three collectives
        k_chars = ctx._empty(int(k_off[-1].item()) + 16, torch.uint8)[:int(k_off[-1].item())]
        ctx._ck(ctx.lib.hmj_gather_spans(ctx.h, chars.data_ptr(), start.data_ptr(), n, k_off.data_ptr(),
                                         k_chars.data_ptr()))
        mark("gather keys")
        rid, r_off, r_chars, _ = exchange_strings(rid_send, offs, k_off, k_chars, self.rank, G, mark=mark,
                                                  key_lens=lens)
        mark("exchange")
        return rid, r_off, r_chars

    def _received(self, side, got):
        """columns of a relation that arrived through the receive buffers"""
        _, nrows, nbytes = got
        rid, lens, chars = self.ctx.xchg_views(side, nrows, nbytes)
        off = torch.zeros(nrows + 1, dtype=torch.int64, device=rid.device)
        if nrows:
            torch.cumsum(lens, 0, dtype=torch.int64, out=off[1:])
        return rid, off, chars

    def _mark(self, what):
        """HMJ_DIST_TRACE=1: device-synchronised wall-clock marks of the string exchange's phases"""
        if not self.trace:
            return
        import time
        torch.cuda.synchronize()
        now = time.perf_counter()
        if what != "start":
            self.trace_log.append((what, (now - self._t_last) * 1e3))
        self._t_last = now

    def join_strings(self, roff_local, rchars_local, soff_local, schars_local, want_pairs=False):
        """String-key join over the ranks (C5's shape).  Each rank passes its block of both
        relations as offsets[n+1] + chars; global row ids are rank-major blocks as in join_u64.
        Returns the number of pairs of this rank's hash range, and with want_pairs the pairs as
        global (build row, probe row) tensors."""
        ctx = self.ctx
        nr, ns = roff_local.numel() - 1, soff_local.numel() - 1
        self.exchange_bytes = 0
        if self.xchg:
            self._rank_barrier()   # nobody writes into a buffer its owner is still joining from
        got_r = self._exchange_side_strings(roff_local, rchars_local, self.rank * nr, side=0)
        got_s = self._exchange_side_strings(soff_local, schars_local, self.rank * ns, side=1)
        if self.xchg:
            self._rank_barrier()   # every rank's sends have landed
            self._mark("start")
        fused = [isinstance(g[0], str) for g in (got_r, got_s)]
        r_id, r_off, r_chars = self._received(0, got_r) if fused[0] else got_r
        s_id, s_off, s_chars = self._received(1, got_s) if fused[1] else got_s
        if all(fused):
            self.xchg_joins += 1
            self._mark("received offsets")
        self.local_rows = (r_id.numel(), s_id.numel())
        # received keys are indexed by arrival position; the join reports positions
        self._mark("start")
        m = ctx.join_strings(r_off, r_chars, s_off, s_chars, prefix_bits=self.bits)
        self._mark("local join")
        if not want_pairs:
            return m
        out_r, out_s = ctx.join_result_tensors(m)
        return m, (r_id[out_r.to(torch.int64)], s_id[out_s.to(torch.int64)])
