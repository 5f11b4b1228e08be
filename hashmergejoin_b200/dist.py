"""Multi-GPU hash-merge join: one process per GPU, hash-range partitioning, ONE exchange.

Replaces the reference's thread-level partitioning (partition_only, /root/reference
partitioned_hash.h:82-124, and the per-thread slices of radix_hash.h:375-388) with the scheme of
SURVEY.md section 8(e): every rank splits its rows of R and S on the top log2(G) hash bits
(K8, hmj_partition_pairs_base -- destination g = hash >> (64-b), exactly the reference's
partition function), the G x G counts are all-gathered, one NCCL all-to-all moves bucket g to
rank g, and each rank then sort-merges its own hash range locally (K2-K6).  Ranges are ordered
by rank, so concatenating the per-rank sorted arrays gives the single-GPU sorted array, and equal
hashes never straddle ranks.  torch.distributed is plumbing only (NCCL all_to_all_single on the
same stream the kernels run on).
"""
import math

import torch
import torch.distributed as dist

from .api import SortedRelation


def plan_exchange(counts_matrix, rank):
    """counts_matrix[src][dst] rows going src->dst.  Returns (send_splits, recv_splits)."""
    send = [int(x) for x in counts_matrix[rank]]
    recv = [int(counts_matrix[src][rank]) for src in range(len(counts_matrix))]
    return send, recv


class DistJoin:
    def __init__(self, ctx, rank, world, rows_total_hint=0):
        assert world & (world - 1) == 0, "hash-range partitioning needs a power-of-two GPU count"
        self.ctx, self.rank, self.world = ctx, rank, world
        self.bits = int(math.log2(world))
        self._pass_pairs = []

    def avg_pass_pairs(self):
        return sum(self._pass_pairs) / max(len(self._pass_pairs), 1)

    def _exchange_side(self, keys, rowid_base):
        """partition local rows by destination, exchange -> (hash, rowid) tensors of my range"""
        ctx, G = self.ctx, self.world
        n = keys.numel()
        ph, pr, offs = ctx.partition_pairs(keys, None, self.bits, rowid_base=rowid_base)
        self._pass_pairs.append(n)
        counts = (offs[1:] - offs[:-1]).contiguous()
        allc = torch.empty(G * G, dtype=torch.int64, device=keys.device)
        dist.all_gather_into_tensor(allc, counts)
        mat = allc.view(G, G).cpu().tolist()
        send, recv = plan_exchange(mat, self.rank)
        nrecv = sum(recv)
        rh = torch.empty(nrecv, dtype=torch.int64, device=keys.device)
        rr = torch.empty(nrecv, dtype=torch.int32, device=keys.device)
        dist.all_to_all_single(rh, ph, recv, send)
        dist.all_to_all_single(rr, pr, recv, send)
        return rh, rr

    def join_u64(self, rkeys_local, skeys_local, want_pairs=False):
        """Returns the number of pairs this rank produced (its hash range)."""
        ctx = self.ctx
        nr, ns = rkeys_local.numel(), skeys_local.numel()
        # global row ids: rank-major blocks (rank r owns rows [r*n, (r+1)*n) of each relation)
        rh, rr = self._exchange_side(rkeys_local, self.rank * nr)
        sh, sr = self._exchange_side(skeys_local, self.rank * ns)
        rh, rr = ctx.sort_pairs(rh, rr, out_hash=rh, out_rowid=rr)
        sh, sr = ctx.sort_pairs(sh, sr, out_hash=sh, out_rowid=sr)
        self._pass_pairs += [rh.numel()] * 8 + [sh.numel()] * 8
        R, S = SortedRelation(rh, rr), SortedRelation(sh, sr)
        if want_pairs:
            out_r, out_s = ctx.merge_join(R, S)
            return out_r.numel(), (out_r, out_s, rh, sh)
        out_r, out_s = ctx.merge_join(R, S)
        return out_r.numel()
