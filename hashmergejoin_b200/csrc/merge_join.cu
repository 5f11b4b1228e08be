// merge_join.cu -- K4-K6: merge-path partitioned join of two sorted (hash,row-id) streams.
//
// Stands in for the sequential merge walk of HashMergeJoin::iterator (/root/reference
// hashjoin.h:70-180: advance the lower hash; on equal hash compare keys) and for its consumer
// loop (hashjoin_bench.cc:131-133).  It computes the TRUE equi-join: every (r,s) whose hashes
// and full keys are equal -- duplicate runs are expanded as the full cross product (the
// reference iterator emits an L-shaped subset for N:M runs and stops at the last element of
// either side; see DESIGN.md "reference defects").
//
//   K4 merge_path_partition_kernel : one diagonal binary search per tile boundary; ties put R
//      first, so for every R element of a tile the lower bound of its hash in S lies inside
//      the tile's S span.
//   K5 join_tile_kernel<COUNT>     : tile = T consecutive elements of the merged order; both
//      spans are staged in shared memory, each R element finds its equal-hash run in S (runs
//      that leave the tile are finished with a galloping search in global memory), candidates
//      are key-verified (strings), per-tile match counts are written.
//      join_scan_kernel            : exclusive scan of the tile counts (+ total).
//   K6 join_tile_kernel<WRITE>     : same walk, writes (build_row, probe_row) pairs at
//      tile_offset + intra-tile exclusive offset.  <AGGREGATE> sums payloads instead.
// Algorithmic bytes: count 8/tuple (12 for strings), write 12/tuple + 8/pair, verification
// gathers per candidate for strings.
#include "common.cuh"

namespace {

constexpr int JOIN_THREADS = 256;
constexpr int JOIN_TILE = 4096;  // merged elements per tile
constexpr size_t JOIN_SMEM = (size_t)JOIN_TILE * 8 + (size_t)JOIN_TILE * 4 + 256;

enum JoinMode { JOIN_COUNT = 0, JOIN_WRITE = 1, JOIN_AGG = 2 };

struct JoinArgs {
  const uint64_t* rh;
  const uint32_t* rr;
  const uint64_t* sh;
  const uint32_t* sr;
  uint32_t nr, ns;
  // string key storage (row-id indexed)
  const uint64_t* roff;
  const uint8_t* rchars;
  const uint64_t* soff;
  const uint8_t* schars;
  // payload columns for aggregate mode (may be null = row id)
  const uint64_t* rvals;
  const uint64_t* svals;
  // partition + tile bookkeeping
  const uint32_t* tile_a;      // [tiles+1] R index at each diagonal
  uint64_t* tile_counts;       // [tiles] matches per tile (count mode writes)
  const uint64_t* tile_offset; // [tiles] exclusive scan (write mode reads)
  uint32_t* out_r;
  uint32_t* out_s;
  uint64_t* agg;               // [2]: pairs, sum
  uint32_t tiles;
};

__device__ __forceinline__ bool keys_equal_str(const uint64_t* __restrict__ aoff,
                                               const uint8_t* __restrict__ achars, uint32_t arow,
                                               const uint64_t* __restrict__ boff,
                                               const uint8_t* __restrict__ bchars, uint32_t brow) {
  const uint64_t a0 = aoff[arow], a1 = aoff[arow + 1];
  const uint64_t b0 = boff[brow], b1 = boff[brow + 1];
  const uint64_t len = a1 - a0;
  if (len != b1 - b0) return false;
  const uint8_t* pa = achars + a0;
  const uint8_t* pb = bchars + b0;
  uint64_t i = 0;
  // 8 bytes at a time when both sides are equally aligned, bytes otherwise
  if (((reinterpret_cast<uintptr_t>(pa) ^ reinterpret_cast<uintptr_t>(pb)) & 7) == 0) {
    while (i < len && (reinterpret_cast<uintptr_t>(pa + i) & 7)) {
      if (pa[i] != pb[i]) return false;
      ++i;
    }
    for (; i + 8 <= len; i += 8)
      if (*reinterpret_cast<const uint64_t*>(pa + i) != *reinterpret_cast<const uint64_t*>(pb + i))
        return false;
  }
  for (; i < len; ++i)
    if (pa[i] != pb[i]) return false;
  return true;
}

// first index in [lo,hi) of sorted a[] with a[idx] >= v   (hi if none)
template <typename P>
__device__ __forceinline__ uint32_t lower_bound_dev(P a, uint32_t lo, uint32_t hi, uint64_t v) {
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    if (a[mid] < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}
// first index in [lo,hi) with a[idx] > v
template <typename P>
__device__ __forceinline__ uint32_t upper_bound_dev(P a, uint32_t lo, uint32_t hi, uint64_t v) {
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    if (a[mid] <= v) lo = mid + 1; else hi = mid;
  }
  return lo;
}
// upper bound of v in global sorted a[from..n) by galloping then bisecting
__device__ __forceinline__ uint32_t gallop_upper(const uint64_t* __restrict__ a, uint32_t from,
                                                 uint32_t n, uint64_t v) {
  if (from >= n || a[from] > v) return from;
  uint32_t lo = from, step = 1;  // invariant: a[lo] <= v
  while (true) {
    const uint64_t nxt = (uint64_t)lo + step;
    if (nxt >= n) break;
    if (a[nxt] > v) return upper_bound_dev(a, lo + 1, (uint32_t)nxt, v);
    lo = (uint32_t)nxt;
    step <<= 1;
  }
  return upper_bound_dev(a, lo + 1, n, v);
}

// K4: R index a_k of diagonal k*T: number of R elements among the first k*T merged elements
// (ties: R before S).
__global__ void merge_path_partition_kernel(const uint64_t* __restrict__ rh, uint32_t nr,
                                            const uint64_t* __restrict__ sh, uint32_t ns,
                                            uint32_t tiles, uint32_t* __restrict__ tile_a) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > tiles) return;
  const uint64_t total = (uint64_t)nr + ns;
  uint64_t diag = (uint64_t)k * JOIN_TILE;
  if (diag > total) diag = total;
  uint64_t lo = diag > ns ? diag - ns : 0;
  uint64_t hi = diag < nr ? diag : nr;
  while (lo < hi) {
    const uint64_t mid = (lo + hi) >> 1;
    // R[mid] precedes S[diag-1-mid] in merged order  <=>  R[mid] <= S[diag-1-mid]
    if (rh[mid] <= sh[diag - 1 - mid]) lo = mid + 1; else hi = mid;
  }
  tile_a[k] = (uint32_t)lo;
}

__device__ __forceinline__ uint64_t block_reduce_u64(uint64_t v, uint64_t* s_tmp) {
  // result valid in thread 0
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) s_tmp[w] = v;
  __syncthreads();
  uint64_t t = 0;
  if (threadIdx.x == 0)
    for (int i = 0; i < JOIN_THREADS / 32; ++i) t += s_tmp[i];
  return t;
}

// exclusive scan of v over the block; *total receives the block sum (all threads)
__device__ __forceinline__ uint64_t block_excl_scan_u64(uint64_t v, uint64_t* s_tmp,
                                                        uint64_t* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint64_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  __syncthreads();
  if (lane == 31) s_tmp[w] = x;
  __syncthreads();
  uint64_t pre = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < JOIN_THREADS / 32; ++i) {
    const uint64_t t = s_tmp[i];
    if (i < w) pre += t;
    tot += t;
  }
  *total = tot;
  return pre + x - v;
}

template <int KIND, int MODE>
__global__ void __launch_bounds__(JOIN_THREADS)
join_tile_kernel(const JoinArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s_h = reinterpret_cast<uint64_t*>(smem_raw);           // [T] R hashes then S hashes
  uint32_t* s_id = reinterpret_cast<uint32_t*>(s_h + JOIN_TILE);   // [T] row ids, same order
  uint64_t* s_tmp = reinterpret_cast<uint64_t*>(s_id + JOIN_TILE); // [8+] scan scratch
  constexpr bool NEED_IDS = (KIND == HMJ_KEY_STRING) || (MODE != JOIN_COUNT);

  const uint32_t tile = blockIdx.x;
  const int tid = threadIdx.x;
  const uint64_t total = (uint64_t)A.nr + A.ns;
  const uint64_t d0 = min((uint64_t)tile * JOIN_TILE, total);
  const uint64_t d1 = min((uint64_t)(tile + 1) * JOIN_TILE, total);
  const uint32_t a0 = A.tile_a[tile], a1 = A.tile_a[tile + 1];
  const uint32_t b0 = (uint32_t)(d0 - a0), b1 = (uint32_t)(d1 - a1);
  const uint32_t na = a1 - a0, nb = b1 - b0;

  for (uint32_t i = tid; i < na; i += JOIN_THREADS) {
    s_h[i] = __ldcs(A.rh + a0 + i);
    if (NEED_IDS) s_id[i] = __ldcs(A.rr + a0 + i);
  }
  for (uint32_t i = tid; i < nb; i += JOIN_THREADS) {
    s_h[na + i] = __ldcs(A.sh + b0 + i);
    if (NEED_IDS) s_id[na + i] = __ldcs(A.sr + b0 + i);
  }
  __syncthreads();
  const uint64_t* s_hs = s_h + na;
  const uint32_t* s_ids = s_id + na;

  uint64_t my_pairs = 0, my_sum = 0;
  uint64_t running = (MODE == JOIN_WRITE) ? A.tile_offset[tile] : 0;

  for (uint32_t base = 0; base < na; base += JOIN_THREADS) {
    const uint32_t li = base + tid;
    const bool active = li < na;
    uint32_t lb = 0, ub = 0, cnt = 0, rrow = 0;
    uint64_t vmask = 0;  // verified-candidate bitmap of the first 64 candidates (strings)
    if (active) {
      const uint64_t h = s_h[li];
      const uint32_t l = lower_bound_dev(s_hs, 0u, nb, h);
      const uint32_t u = upper_bound_dev(s_hs, l, nb, h);
      lb = b0 + l;
      ub = (u == nb) ? gallop_upper(A.sh, b1, A.ns, h) : b0 + u;
      if (NEED_IDS) rrow = s_id[li];
      if (KIND == HMJ_KEY_U64) {
        cnt = ub - lb;
      } else {
        for (uint32_t j = lb; j < ub; ++j) {
          const uint32_t srow = (j < b1) ? s_ids[j - b0] : A.sr[j];
          if (keys_equal_str(A.roff, A.rchars, rrow, A.soff, A.schars, srow)) {
            if (j - lb < 64) vmask |= 1ull << (j - lb);
            ++cnt;
          }
        }
      }
    }
    if (MODE == JOIN_COUNT) {
      my_pairs += cnt;
    } else {
      uint64_t off = 0;
      if (MODE == JOIN_WRITE) {
        uint64_t round_total;
        off = running + block_excl_scan_u64(cnt, s_tmp, &round_total);
        running += round_total;
      }
      if (active && cnt) {
        uint64_t rv = 0;
        if (MODE == JOIN_AGG) rv = A.rvals ? A.rvals[rrow] : (uint64_t)rrow;
        for (uint32_t j = lb; j < ub; ++j) {
          const uint32_t srow = (j < b1) ? s_ids[j - b0] : A.sr[j];
          if (KIND == HMJ_KEY_STRING) {
            const bool ok = (j - lb < 64)
                                ? ((vmask >> (j - lb)) & 1)
                                : keys_equal_str(A.roff, A.rchars, rrow, A.soff, A.schars, srow);
            if (!ok) continue;
          }
          if (MODE == JOIN_WRITE) {
            A.out_r[off] = rrow;
            A.out_s[off] = srow;
            ++off;
          } else {
            my_sum += rv + (A.svals ? A.svals[srow] : (uint64_t)srow);
          }
        }
        my_pairs += cnt;
      }
    }
  }
  if (MODE == JOIN_COUNT) {
    const uint64_t t = block_reduce_u64(my_pairs, s_tmp);
    if (tid == 0) A.tile_counts[tile] = t;
  } else if (MODE == JOIN_AGG) {
    const uint64_t p = block_reduce_u64(my_pairs, s_tmp);
    const uint64_t s = block_reduce_u64(my_sum, s_tmp);
    if (tid == 0) {
      if (p) atomicAdd(reinterpret_cast<unsigned long long*>(A.agg), (unsigned long long)p);
      if (s) atomicAdd(reinterpret_cast<unsigned long long*>(A.agg + 1), (unsigned long long)s);
    }
  }
}

// exclusive scan of tile counts by one block; total -> out_total
__global__ void __launch_bounds__(1024)
join_scan_kernel(const uint64_t* __restrict__ counts, uint32_t tiles,
                 uint64_t* __restrict__ offsets, uint64_t* __restrict__ out_total) {
  __shared__ uint64_t s_w[32];
  __shared__ uint64_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < tiles; base += 1024) {
    const uint32_t i = base + tid;
    const uint64_t v = i < tiles ? counts[i] : 0;
    uint64_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_w[w] = x;
    __syncthreads();
    uint64_t pre = s_carry;
    for (int k = 0; k < w; ++k) pre += s_w[k];
    if (i < tiles) offsets[i] = pre + x - v;
    __syncthreads();
    if (tid == 1023) s_carry = pre + x;
    __syncthreads();
  }
  if (tid == 0) *out_total = s_carry;
}

}  // namespace

// ---------------------------------------------------------------- host side
namespace {

int check_relation(hmj_ctx* ctx, const hmj_relation* r) {
  HMJ_ARG(ctx, r != nullptr);
  HMJ_ARG(ctx, r->n < (size_t(1) << 32) - (size_t(1) << 16));
  HMJ_ARG(ctx, r->n == 0 || (r->d_hash && r->d_rowid));
  HMJ_ARG(ctx, r->key_kind == HMJ_KEY_U64 || r->key_kind == HMJ_KEY_STRING);
  if (r->key_kind == HMJ_KEY_STRING) HMJ_ARG(ctx, r->n == 0 || (r->d_offsets && r->d_chars));
  return HMJ_OK;
}

struct JoinPlan {
  JoinArgs args;
  int kind;
};

// K4: partition.  Fills the plan; tile_counts / tile_offset live in the workspace.
int join_plan(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s, JoinPlan* plan) {
  HMJ_TRY(check_relation(ctx, r));
  HMJ_TRY(check_relation(ctx, s));
  HMJ_ARG(ctx, r->key_kind == s->key_kind);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  JoinArgs& A = plan->args;
  memset(&A, 0, sizeof A);
  plan->kind = r->key_kind;
  A.rh = r->d_hash; A.rr = r->d_rowid; A.nr = (uint32_t)r->n;
  A.sh = s->d_hash; A.sr = s->d_rowid; A.ns = (uint32_t)s->n;
  A.roff = r->d_offsets; A.rchars = r->d_chars;
  A.soff = s->d_offsets; A.schars = s->d_chars;
  const uint64_t total = (uint64_t)r->n + s->n;
  A.tiles = (uint32_t)hmj_div_up(total, JOIN_TILE);
  uint32_t* tile_a = nullptr;
  // layout of WS_JOIN_TILES: [tiles+1] u32 tile_a | [tiles] u64 counts | [tiles] u64 offsets
  const size_t a_bytes = (((size_t)A.tiles + 1) * 4 + 15) & ~size_t(15);
  void* base = nullptr;
  HMJ_TRY(hmj_ws_get(ctx, WS_JOIN_TILES, a_bytes + (size_t)A.tiles * 16 + 16, &base));
  tile_a = reinterpret_cast<uint32_t*>(base);
  A.tile_a = tile_a;
  A.tile_counts = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(base) + a_bytes);
  A.tile_offset = A.tile_counts + A.tiles;
  uint64_t* scal = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_SCALARS, 8, &scal));
  A.agg = scal;  // [0],[1] aggregate; [2] scan total
  if (A.tiles == 0) return HMJ_OK;
  const uint32_t nthreads = A.tiles + 1;
  merge_path_partition_kernel<<<(nthreads + 255) / 256, 256, 0, ctx->stream>>>(
      A.rh, A.nr, A.sh, A.ns, A.tiles, tile_a);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

template <int MODE>
int launch_tiles(hmj_ctx* ctx, const JoinPlan& plan) {
  if (plan.args.tiles == 0) return HMJ_OK;
  if (plan.kind == HMJ_KEY_U64) {
    HMJ_CUDA(ctx, cudaFuncSetAttribute(join_tile_kernel<HMJ_KEY_U64, MODE>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)JOIN_SMEM));
    join_tile_kernel<HMJ_KEY_U64, MODE><<<plan.args.tiles, JOIN_THREADS, JOIN_SMEM, ctx->stream>>>(plan.args);
  } else {
    HMJ_CUDA(ctx, cudaFuncSetAttribute(join_tile_kernel<HMJ_KEY_STRING, MODE>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)JOIN_SMEM));
    join_tile_kernel<HMJ_KEY_STRING, MODE><<<plan.args.tiles, JOIN_THREADS, JOIN_SMEM, ctx->stream>>>(plan.args);
  }
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

// K5 + scan: exact pair count on the host (synchronises the stream)
int join_count(hmj_ctx* ctx, JoinPlan* plan, uint64_t* h_total) {
  *h_total = 0;
  if (plan->args.tiles == 0) return HMJ_OK;
  HMJ_TRY(launch_tiles<JOIN_COUNT>(ctx, *plan));
  uint64_t* d_total = plan->args.agg + 2;
  join_scan_kernel<<<1, 1024, 0, ctx->stream>>>(plan->args.tile_counts, plan->args.tiles,
                                               const_cast<uint64_t*>(plan->args.tile_offset), d_total);
  HMJ_LAUNCHED(ctx);
  uint64_t* h = reinterpret_cast<uint64_t*>(ctx->pinned);
  HMJ_CUDA(ctx, cudaMemcpyAsync(h, d_total, 8, cudaMemcpyDeviceToHost, ctx->stream));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *h_total = h[0];
  return HMJ_OK;
}

}  // namespace

extern "C" int hmj_merge_join_count(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s,
                                    uint64_t* h_num_pairs) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, h_num_pairs != nullptr);
  JoinPlan plan;
  HMJ_TRY(join_plan(ctx, r, s, &plan));
  return join_count(ctx, &plan, h_num_pairs);
}

extern "C" int hmj_merge_join_write(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s,
                                    uint32_t* d_out_r, uint32_t* d_out_s, size_t cap,
                                    uint64_t* h_num_pairs) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, h_num_pairs != nullptr);
  JoinPlan plan;
  HMJ_TRY(join_plan(ctx, r, s, &plan));
  HMJ_TRY(join_count(ctx, &plan, h_num_pairs));
  if (*h_num_pairs > cap)
    return hmj_fail(ctx, HMJ_ECAPACITY, "output capacity too small", __FILE__, __LINE__);
  if (*h_num_pairs == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_out_r && d_out_s);
  plan.args.out_r = d_out_r;
  plan.args.out_s = d_out_s;
  return launch_tiles<JOIN_WRITE>(ctx, plan);
}

extern "C" int hmj_merge_join_aggregate(hmj_ctx* ctx, const hmj_relation* r,
                                        const hmj_relation* s, const uint64_t* d_rvals,
                                        const uint64_t* d_svals, uint64_t* h_num_pairs,
                                        uint64_t* h_sum) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, h_num_pairs && h_sum);
  JoinPlan plan;
  HMJ_TRY(join_plan(ctx, r, s, &plan));
  *h_num_pairs = 0;
  *h_sum = 0;
  if (plan.args.tiles == 0) return HMJ_OK;
  plan.args.rvals = d_rvals;
  plan.args.svals = d_svals;
  HMJ_CUDA(ctx, cudaMemsetAsync(plan.args.agg, 0, 16, ctx->stream));
  HMJ_TRY(launch_tiles<JOIN_AGG>(ctx, plan));
  uint64_t* h = reinterpret_cast<uint64_t*>(ctx->pinned);
  HMJ_CUDA(ctx, cudaMemcpyAsync(h, plan.args.agg, 16, cudaMemcpyDeviceToHost, ctx->stream));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *h_num_pairs = h[0];
  *h_sum = h[1];
  return HMJ_OK;
}
