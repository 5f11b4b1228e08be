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
#include "join_device.cuh"

namespace {

constexpr int JOIN_THREADS = 256;
#ifndef HMJ_JOIN_TILE
#define HMJ_JOIN_TILE 4096
#endif
constexpr int JOIN_TILE = HMJ_JOIN_TILE;  // merged elements per tile
constexpr int JOIN_VT = JOIN_TILE / JOIN_THREADS;  // merged elements per thread in the tile merge
// hashes are stored padded (one extra slot per 8) so that threads walking their own chunk of the
// merged order (stride ~8 per array) hit different banks
__host__ __device__ constexpr uint32_t jpad(uint32_t i) { return i + (i >> 3); }
constexpr size_t JOIN_HASH_SLOTS = JOIN_TILE + JOIN_TILE / 8 + 2;
constexpr size_t JOIN_SMEM_BASE = JOIN_HASH_SLOTS * 8 + (size_t)JOIN_TILE * 4 + 8192;
constexpr size_t JOIN_SMEM_STR = JOIN_SMEM_BASE + (size_t)JOIN_TILE * 4;  // + verified counts

enum JoinMode { JOIN_COUNT = 0, JOIN_WRITE = 1, JOIN_AGG = 2, JOIN_FUSED = 3 };

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
  uint64_t* agg;               // [0],[1]: pairs, sum (aggregate); [2]: total pairs
  uint64_t* tile_desc;         // [tiles] look-back descriptors (fused mode)
  uint32_t* tile_counter;      // dynamic tile ids (fused mode)
  uint64_t cap;                // capacity of out_r/out_s in pairs (fused mode)
  uint32_t tiles;
  // Fused mode, string keys: count every hash-equal candidate as a match and verify the keys only
  // once, where the pair is written (a candidate's key bytes are a random gather; verifying in the
  // count phase AND in the write phase made the join of C5 gather every probe key twice).  A
  // candidate that fails the key check -- two different keys with one 64-bit hash -- raises
  // *collision and the host re-runs the exact two-phase join.
  int optimistic;
  uint32_t* collision;
};

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

// ---- cooperative handling of long runs -------------------------------------------------
// A probe (R) element whose equal-hash run in S is longer than JOIN_BIG is not walked by its
// own thread: the owner parks its parameters in shared memory and the whole CTA walks the run
// with coalesced accesses (counting verified candidates, or emitting pairs).
constexpr int JOIN_BIG = 64;

struct JoinShared {  // lives behind the tile arrays in dynamic shared memory
  uint64_t tmp[JOIN_THREADS / 32 + 2];
  uint64_t p_off[JOIN_THREADS];
  uint32_t p_rrow[JOIN_THREADS], p_lb[JOIN_THREADS], p_n[JOIN_THREADS];  // p_n: in candidates, out verified
  uint32_t q[JOIN_THREADS];
  uint64_t wt[JOIN_TILE / 32];  // fused mode: match totals per (round, warp), then their scan
  uint32_t qn;
  uint32_t tile;
  uint64_t base;
};

template <int KIND>
__device__ __forceinline__ bool cand_ok(const JoinArgs& A, uint32_t rrow, uint32_t srow) {
  if (KIND == HMJ_KEY_U64) return true;
  const bool ok = keys_equal_str(A.roff, A.rchars, rrow, A.soff, A.schars, srow);
  if (!ok && A.optimistic) *reinterpret_cast<volatile uint32_t*>(A.collision) = 1u;
  return ok;
}

// all threads: verified candidates of every queued run -> p_res[owner]
template <int KIND>
__device__ void coop_count(const JoinArgs& A, JoinShared* sh) {
  __syncthreads();
  const uint32_t qn = sh->qn;
  for (uint32_t e = 0; e < qn; ++e) {
    const uint32_t o = sh->q[e];
    const uint32_t rrow = sh->p_rrow[o], lb = sh->p_lb[o], n = sh->p_n[o];
    uint64_t c = 0;
    for (uint32_t t = threadIdx.x; t < n; t += JOIN_THREADS)
      c += cand_ok<KIND>(A, rrow, A.sr[lb + t]) ? 1 : 0;
    const uint64_t tot = block_reduce_u64(c, sh->tmp);
    if (threadIdx.x == 0) sh->p_n[o] = (uint32_t)tot;  // every thread has read p_n[o] (reduce syncs)
  }
  __syncthreads();
  if (threadIdx.x == 0) sh->qn = 0;
  __syncthreads();
}

// all threads: emit (or aggregate) the pairs of every queued run
template <int KIND, int MODE>
__device__ void coop_emit(const JoinArgs& A, JoinShared* sh, bool do_write, uint64_t* my_sum) {
  __syncthreads();
  const uint32_t qn = sh->qn;
  for (uint32_t e = 0; e < qn; ++e) {
    const uint32_t o = sh->q[e];
    const uint32_t rrow = sh->p_rrow[o], lb = sh->p_lb[o], n = sh->p_n[o];
    uint64_t off = sh->p_off[o];
    const uint64_t rv = (MODE == JOIN_AGG) ? (A.rvals ? A.rvals[rrow] : (uint64_t)rrow) : 0;
    for (uint32_t c0 = 0; c0 < n; c0 += JOIN_THREADS) {
      const uint32_t t = c0 + threadIdx.x;
      uint32_t srow = 0;
      bool ok = false;
      if (t < n) {
        srow = A.sr[lb + t];
        ok = cand_ok<KIND>(A, rrow, srow);
      }
      if (MODE == JOIN_AGG) {
        if (ok) *my_sum += rv + (A.svals ? A.svals[srow] : (uint64_t)srow);
      } else if (KIND == HMJ_KEY_U64) {
        if (ok && do_write) {
          A.out_r[off + t] = rrow;
          A.out_s[off + t] = srow;
        }
      } else {
        uint64_t tot;
        const uint64_t pre = block_excl_scan_u64(ok ? 1 : 0, sh->tmp, &tot);
        if (ok && do_write) {
          A.out_r[off + pre] = rrow;
          A.out_s[off + pre] = srow;
        }
        off += tot;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) sh->qn = 0;
  __syncthreads();
}

static_assert(sizeof(JoinShared) <= 8192, "JOIN_SMEM reserves 8 KB for JoinShared");
constexpr uint64_t JD_PARTIAL = 1ull << 62;
constexpr uint64_t JD_INCLUSIVE = 2ull << 62;
constexpr uint64_t JD_VALUE = (1ull << 62) - 1;

__device__ __forceinline__ void jd_store(uint64_t* p, uint64_t v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t jd_load(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// One tile of the merged order.  MODE:
//   JOIN_COUNT : per-tile match counts (tile_counts) only
//   JOIN_WRITE : offsets known (tile_offset, after join_scan_kernel): write the pairs
//   JOIN_AGG   : count + sum of payloads, nothing materialised
//   JOIN_FUSED : single pass -- count the tile, obtain its output offset by a decoupled look-back
//                over per-tile descriptors, write the pairs if they fit in `cap`; tile_counts are
//                published too, so an overflowing join can finish with scan + JOIN_WRITE.
template <int KIND, int MODE>
__global__ void __launch_bounds__(JOIN_THREADS, 2)
join_tile_kernel(const JoinArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s_h = reinterpret_cast<uint64_t*>(smem_raw);               // R hashes then S hashes, padded
  // [T] per R element: lower bound in the S span; fused mode later reuses the slot for the
  // element's output offset inside its (round, warp) group
  uint32_t* s_lb = reinterpret_cast<uint32_t*>(s_h + JOIN_HASH_SLOTS);
  JoinShared* sh = reinterpret_cast<JoinShared*>(s_lb + JOIN_TILE);
  uint32_t* s_cnt = reinterpret_cast<uint32_t*>(reinterpret_cast<unsigned char*>(sh) + 8192);  // strings
  constexpr bool NEED_IDS = (KIND == HMJ_KEY_STRING) || (MODE != JOIN_COUNT);
  const int tid = threadIdx.x;

  if (tid == 0) {
    sh->qn = 0;
    sh->tile = (MODE == JOIN_FUSED) ? atomicAdd(A.tile_counter, 1u) : blockIdx.x;
  }
  __syncthreads();
  const uint32_t tile = sh->tile;
  const uint64_t total = (uint64_t)A.nr + A.ns;
  const uint64_t d0 = min((uint64_t)tile * JOIN_TILE, total);
  const uint64_t d1 = min((uint64_t)(tile + 1) * JOIN_TILE, total);
  const uint32_t a0 = A.tile_a[tile], a1 = A.tile_a[tile + 1];
  const uint32_t b0 = (uint32_t)(d0 - a0), b1 = (uint32_t)(d1 - a1);
  const uint32_t na = a1 - a0, nb = b1 - b0;

  // R hashes at padded slots [0,na), S hashes at padded slots of (na + i)
  // (row ids are not staged: they are read from global memory where pairs are emitted --
  // consecutive threads read consecutive entries)
  for (uint32_t i = tid; i < na; i += JOIN_THREADS) s_h[jpad(i)] = __ldcs(A.rh + a0 + i);
  for (uint32_t i = tid; i < nb; i += JOIN_THREADS) s_h[jpad(na + i)] = __ldcs(A.sh + b0 + i);
  __syncthreads();
  auto RH = [&](uint32_t i) -> uint64_t { return s_h[jpad(i)]; };
  auto SH = [&](uint32_t i) -> uint64_t { return s_h[jpad(na + i)]; };

  // ---- tile merge: every thread walks JOIN_VT consecutive elements of the merged order (ties:
  // R first) and records, for each R element it passes, how many S elements precede it = the
  // lower bound of its hash in the S span.
  {
    const uint32_t nt = na + nb;
    const uint32_t diag = min((uint32_t)tid * JOIN_VT, nt);
    uint32_t lo = diag > nb ? diag - nb : 0, hi = min(diag, na);
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      if (RH(mid) <= SH(diag - 1 - mid)) lo = mid + 1; else hi = mid;
    }
    uint32_t a = lo, b = diag - lo;
    uint64_t ra = a < na ? RH(a) : 0, sb = b < nb ? SH(b) : 0;
#pragma unroll 4
    for (int i = 0; i < JOIN_VT; ++i) {
      if (a + b >= nt) break;
      const bool take_r = (b >= nb) || (a < na && ra <= sb);
      if (take_r) {
        s_lb[a] = b;
        ++a;
        if (a < na) ra = RH(a);
      } else {
        ++b;
        if (b < nb) sb = SH(b);
      }
    }
  }
  __syncthreads();

  uint64_t my_pairs = 0, my_sum = 0;
  uint64_t running = (MODE == JOIN_WRITE) ? A.tile_offset[tile] : 0;
  bool do_write = (MODE == JOIN_WRITE);
  const uint32_t rounds = (na + JOIN_THREADS - 1) / JOIN_THREADS;

  // emits the pairs of one round: per-thread (rrow, lb, ncand, cnt), offsets from a block scan
  auto emit_round = [&](bool active, uint32_t rrow, uint32_t lb, uint32_t ncand, uint32_t cnt,
                        uint64_t vmask, bool have_mask) {
    uint64_t off = 0;
    if (MODE != JOIN_AGG) {
      uint64_t round_total;
      off = running + block_excl_scan_u64(cnt, sh->tmp, &round_total);
      running += round_total;
    }
    const bool big = active && ncand > (uint32_t)JOIN_BIG;
    if (big) {
      sh->p_rrow[tid] = rrow; sh->p_lb[tid] = lb; sh->p_n[tid] = ncand; sh->p_off[tid] = off;
      sh->q[atomicAdd(&sh->qn, 1u)] = tid;
    } else if (active && cnt && (do_write || MODE == JOIN_AGG)) {
      const uint64_t rv = (MODE == JOIN_AGG) ? (A.rvals ? A.rvals[rrow] : (uint64_t)rrow) : 0;
      for (uint32_t t = 0; t < ncand; ++t) {
        const uint32_t j = lb + t;
        const uint32_t srow = A.sr[j];
        if (KIND == HMJ_KEY_STRING) {
          const bool ok = (have_mask && t < 64) ? ((vmask >> t) & 1) : cand_ok<KIND>(A, rrow, srow);
          if (!ok) continue;
        }
        if (MODE == JOIN_AGG) {
          my_sum += rv + (A.svals ? A.svals[srow] : (uint64_t)srow);
        } else {
          A.out_r[off] = rrow;
          A.out_s[off] = srow;
          ++off;
        }
      }
    }
    if (__syncthreads_or(big)) coop_emit<KIND, MODE>(A, sh, do_write, &my_sum);
  };

  bool wide = false, any_big = false;  // fused: an offset needs 64 bits / a run is walked by the CTA
  // ---- phase A: equal-hash run of every R element, verified match count
  for (uint32_t r = 0; r < rounds; ++r) {
    const uint32_t li = r * JOIN_THREADS + tid;
    const bool active = li < na;
    uint32_t lb = 0, ncand = 0, cnt = 0, rrow = 0;
    uint64_t vmask = 0;  // verified-candidate bitmap of the first 64 candidates (strings)
    bool big = false;
    if (active) {
      const uint64_t h = RH(li);
      const uint32_t l = s_lb[li];
      // upper bound: short runs by stepping, longer ones by bisection in the S span
      uint32_t u = l;
      while (u < nb && u < l + 4 && SH(u) == h) ++u;
      if (u == l + 4 && u < nb && SH(u) == h) {
        uint32_t lo2 = u + 1, hi2 = nb;
        while (lo2 < hi2) {
          const uint32_t mid = lo2 + ((hi2 - lo2) >> 1);
          if (SH(mid) <= h) lo2 = mid + 1; else hi2 = mid;
        }
        u = lo2;
      }
      lb = b0 + l;
      const uint32_t ub = (u == nb) ? gallop_upper(A.sh, b1, A.ns, h) : b0 + u;
      ncand = ub - lb;
      if (NEED_IDS) rrow = A.rr[a0 + li];
      if (KIND == HMJ_KEY_U64 || (MODE == JOIN_FUSED && A.optimistic)) {
        cnt = ncand;
      } else if (ncand > (uint32_t)JOIN_BIG) {
        big = true;
        sh->p_rrow[tid] = rrow; sh->p_lb[tid] = lb; sh->p_n[tid] = ncand;
        sh->q[atomicAdd(&sh->qn, 1u)] = tid;
      } else {
        for (uint32_t t = 0; t < ncand; ++t) {
          const uint32_t j = lb + t;
          const uint32_t srow = A.sr[j];
          if (cand_ok<KIND>(A, rrow, srow)) {
            vmask |= 1ull << t;
            ++cnt;
          }
        }
      }
    }
    if (KIND == HMJ_KEY_STRING) {
      if (__syncthreads_or(big)) {
        coop_count<KIND>(A, sh);
        if (big) cnt = sh->p_n[tid];
      }
    }
    my_pairs += cnt;
    if (MODE == JOIN_WRITE || MODE == JOIN_AGG) {
      emit_round(active, rrow, lb, ncand, cnt, vmask, true);
    } else if (MODE == JOIN_FUSED) {
      // offsets: inclusive scan of the match counts inside the warp now, (round, warp) totals
      // are scanned once per tile below -- no block-wide scan per round
      uint64_t x = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
        if ((tid & 31) >= o) x += y;
      }
      const uint64_t wexcl = x - cnt;
      if ((tid & 31) == 31) sh->wt[r * (JOIN_THREADS / 32) + (tid >> 5)] = x;
      if (active) {
        // this thread alone reads R hash li from here on: reuse its slot for (lb, ncand)
        s_h[jpad(li)] = (uint64_t)lb | ((uint64_t)ncand << 32);
        s_lb[li] = (uint32_t)wexcl;
        if (KIND == HMJ_KEY_STRING) s_cnt[li] = cnt;
        if (wexcl >> 32) wide = true;
        if (ncand > (uint32_t)JOIN_BIG) any_big = true;
      }
    }
  }

  if (MODE == JOIN_COUNT) {
    const uint64_t t = block_reduce_u64(my_pairs, sh->tmp);
    if (tid == 0) A.tile_counts[tile] = t;
  } else if (MODE == JOIN_FUSED) {
    const int flags = __syncthreads_or((wide ? 1 : 0) | (any_big ? 2 : 0));
    if (tid < 32) {
      // exclusive scan of the (round, warp) totals, in element order
      constexpr int PER_LANE = JOIN_TILE / 32 / 32;
      static_assert(PER_LANE >= 1, "tile too small for the warp scan layout");
      const uint32_t groups = rounds * (JOIN_THREADS / 32);
      uint64_t v[PER_LANE], sum = 0;
#pragma unroll
      for (int i = 0; i < PER_LANE; ++i) {
        const uint32_t g = tid * PER_LANE + i;
        v[i] = g < groups ? sh->wt[g] : 0;
        sum += v[i];
      }
      uint64_t x = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (tid >= o) x += y;
      }
      const uint64_t mine = __shfl_sync(0xffffffffu, x, 31);  // matches of the whole tile
      uint64_t run = x - sum;
#pragma unroll
      for (int i = 0; i < PER_LANE; ++i) {
        const uint32_t g = tid * PER_LANE + i;
        if (g < groups) sh->wt[g] = run;
        run += v[i];
      }
      if (tid == 0) {
        A.tile_counts[tile] = mine;
        sh->p_off[0] = mine;
        if (tile != 0) jd_store(A.tile_desc + tile, JD_PARTIAL | mine);
      }
      // decoupled look-back, one warp: 32 predecessor descriptors per round trip
      uint64_t excl = 0;
      int64_t first_idx = (int64_t)tile - 1;
      while (first_idx >= 0) {
        const int64_t idx = first_idx - tid;
        uint64_t d = JD_INCLUSIVE;  // before tile 0: an inclusive prefix of zero
        if (idx >= 0) {
          d = jd_load(A.tile_desc + idx);
          while ((d >> 62) == 0) d = jd_load(A.tile_desc + idx);
        }
        const uint32_t incl = __ballot_sync(0xffffffffu, (d >> 62) == 2);
        const int stop = incl ? (__ffs(incl) - 1) : 31;
        uint64_t c = (tid <= stop) ? (d & JD_VALUE) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        excl += c;
        if (incl) break;
        first_idx -= 32;
      }
      if (tid == 0) {
        jd_store(A.tile_desc + tile, JD_INCLUSIVE | (excl + mine));
        if (tile == A.tiles - 1) A.agg[2] = excl + mine;
        sh->base = excl;
      }
    }
    __syncthreads();
    const uint64_t tile_base = sh->base;
    const uint64_t tile_total = sh->p_off[0];
    do_write = tile_base + tile_total <= A.cap;
    if (tile_total != 0 && do_write) {
      // ---- phase B: emit with the offsets now known
      if (flags == 0) {
        // common case: short runs only, 32-bit offsets inside a (round, warp) group.  Rounds
        // are taken four at a time so that the row-id loads of a group are all in flight
        // before the first store waits on one of them.
        constexpr int G = 4;
        for (uint32_t r0 = 0; r0 < rounds; r0 += G) {
          uint32_t lb[G], nc[G], rrow[G], srow[G];
          uint64_t off[G];
#pragma unroll
          for (int g = 0; g < G; ++g) {
            const uint32_t r = r0 + g, li = r * JOIN_THREADS + tid;
            nc[g] = 0;
            if (r < rounds && li < na) {
              const uint64_t w = s_h[jpad(li)];
              lb[g] = (uint32_t)w;
              nc[g] = (uint32_t)(w >> 32);
              if (nc[g]) {
                off[g] = tile_base + sh->wt[r * (JOIN_THREADS / 32) + (tid >> 5)] + s_lb[li];
                rrow[g] = A.rr[a0 + li];
                srow[g] = A.sr[lb[g]];
              }
            }
          }
          if (KIND == HMJ_KEY_STRING) {
            // String keys: the first two candidates of two elements at a time, with the loads of
            // all six keys staged (offsets, then bytes, then the comparison) instead of one
            // dependent chain per candidate (join_device.cuh).  Further candidates, empty keys
            // and keys longer than 32 bytes take the one-by-one check.
#pragma unroll
            for (int h0 = 0; h0 < G; h0 += 2) {
              uint32_t sr1[2];
              hmj_key_span sa[2], sb0[2], sb1[2];
              bool fast[2];
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int g = h0 + h;
                sr1[h] = 0;
                fast[h] = false;
                if (nc[g]) {
                  sa[h] = key_span_of(A.roff, rrow[g]);
                  sb0[h] = key_span_of(A.soff, srow[g]);
                  sr1[h] = nc[g] > 1 ? A.sr[lb[g] + 1] : srow[g];
                }
              }
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int g = h0 + h;
                if (nc[g]) {
                  sb1[h] = key_span_of(A.soff, sr1[h]);
                  fast[h] = sa[h].len >= 1 && sa[h].len <= 32;
                }
              }
              hmj_key_words ka[2], kb0[2], kb1[2];
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                if (fast[h]) {
                  key_words_load(A.rchars, sa[h], &ka[h]);
                  // (a candidate of another length cannot match: load nothing for it)
                  if (sb0[h].len == sa[h].len) key_words_load(A.schars, sb0[h], &kb0[h]);
                  if (sb1[h].len == sa[h].len) key_words_load(A.schars, sb1[h], &kb1[h]);
                }
              }
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int g = h0 + h;
                if (nc[g] == 0) continue;
                uint64_t o = off[g];
                for (uint32_t t = 0; t < nc[g]; ++t) {
                  const uint32_t sr = t == 0 ? srow[g] : t == 1 ? sr1[h] : A.sr[lb[g] + t];
                  bool ok;
                  if (fast[h] && t < 2) {
                    const hmj_key_span& sb = t == 0 ? sb0[h] : sb1[h];
                    ok = sb.len == sa[h].len &&
                         key_words_equal(ka[h], t == 0 ? kb0[h] : kb1[h], (uint32_t)sa[h].len);
                    if (!ok && A.optimistic) *reinterpret_cast<volatile uint32_t*>(A.collision) = 1u;
                  } else {
                    ok = cand_ok<KIND>(A, rrow[g], sr);
                  }
                  if (!ok) continue;
                  A.out_r[o] = rrow[g];
                  A.out_s[o] = sr;
                  ++o;
                }
              }
            }
          } else {
#pragma unroll
            for (int g = 0; g < G; ++g) {
              if (nc[g] == 0) continue;
              uint64_t o = off[g];
              for (uint32_t t = 0; t < nc[g]; ++t) {
                const uint32_t sr = t == 0 ? srow[g] : A.sr[lb[g] + t];
                A.out_r[o] = rrow[g];
                A.out_s[o] = sr;
                ++o;
              }
            }
          }
        }
      } else {
        // long runs (walked by the whole CTA) or 64-bit offsets: block scan per round
        running = tile_base;
        for (uint32_t r = 0; r < rounds; ++r) {
          const uint32_t li = r * JOIN_THREADS + tid;
          const bool active = li < na;
          uint32_t lb = 0, ncand = 0, cnt = 0, rrow = 0;
          if (active) {
            const uint64_t w = s_h[jpad(li)];
            lb = (uint32_t)w;
            ncand = (uint32_t)(w >> 32);
            cnt = (KIND == HMJ_KEY_STRING) ? s_cnt[li] : ncand;
            rrow = A.rr[a0 + li];
          }
          emit_round(active, rrow, lb, ncand, cnt, 0, false);
        }
      }
    }
  } else if (MODE == JOIN_AGG) {
    const uint64_t p = block_reduce_u64(my_pairs, sh->tmp);
    const uint64_t s = block_reduce_u64(my_sum, sh->tmp);
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
  // layout of WS_JOIN_TILES: [tiles+1] u32 tile_a | [tiles] u64 counts | [tiles] u64 offsets |
  // [tiles] u64 look-back descriptors | u32 tile counter
  const size_t a_bytes = (((size_t)A.tiles + 1) * 4 + 15) & ~size_t(15);
  void* base = nullptr;
  HMJ_TRY(hmj_ws_get(ctx, WS_JOIN_TILES, a_bytes + (size_t)A.tiles * 24 + 32, &base));
  tile_a = reinterpret_cast<uint32_t*>(base);
  A.tile_a = tile_a;
  A.tile_counts = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(base) + a_bytes);
  A.tile_offset = A.tile_counts + A.tiles;
  A.tile_desc = const_cast<uint64_t*>(A.tile_offset) + A.tiles;
  A.tile_counter = reinterpret_cast<uint32_t*>(A.tile_desc + A.tiles);
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
  cudaEvent_t ev;  // algorithmic bytes: 12 per input tuple here, + 8 per pair once the count is known
  HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_JOIN, ((double)plan.args.nr + plan.args.ns) * (MODE == JOIN_COUNT ? 8.0 : 12.0), &ev));
  if (plan.kind == HMJ_KEY_U64) {
    HMJ_CUDA(ctx, cudaFuncSetAttribute(join_tile_kernel<HMJ_KEY_U64, MODE>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)JOIN_SMEM_BASE));
    join_tile_kernel<HMJ_KEY_U64, MODE><<<plan.args.tiles, JOIN_THREADS, JOIN_SMEM_BASE, ctx->stream>>>(plan.args);
  } else {
    HMJ_CUDA(ctx, cudaFuncSetAttribute(join_tile_kernel<HMJ_KEY_STRING, MODE>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)JOIN_SMEM_STR));
    join_tile_kernel<HMJ_KEY_STRING, MODE><<<plan.args.tiles, JOIN_THREADS, JOIN_SMEM_STR, ctx->stream>>>(plan.args);
  }
  HMJ_LAUNCHED(ctx);
  hmj_prof_end(ctx, ev);
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

// Fused count + write in one pass over the tiles; pairs are written only while they fit in
// `cap`.  Returns the exact total; when it exceeds cap nothing useful was written and the
// caller re-runs with a larger buffer (tile_counts are valid: scan + JOIN_WRITE would do too).
int join_fused(hmj_ctx* ctx, JoinPlan* plan, uint32_t* d_out_r, uint32_t* d_out_s, uint64_t cap,
               uint64_t* h_total) {
  *h_total = 0;
  JoinArgs& A = plan->args;
  if (A.tiles == 0) return HMJ_OK;
  A.out_r = d_out_r;
  A.out_s = d_out_s;
  A.cap = cap;
  static const bool allow_optimistic = [] {
    const char* e = getenv("HMJ_JOIN_OPTIMISTIC");
    return !(e && atoi(e) == 0);
  }();
  uint64_t* h = reinterpret_cast<uint64_t*>(ctx->pinned);
  A.collision = A.tile_counter + 2;  // inside the zeroed tail of the descriptors
  for (int attempt = 0; attempt < 2; ++attempt) {
    A.optimistic = (attempt == 0 && allow_optimistic && plan->kind == HMJ_KEY_STRING) ? 1 : 0;
    HMJ_TRY(hmj_fill_async(ctx, A.tile_desc, 0, (size_t)A.tiles * 8 + 16, ctx->stream));
    HMJ_TRY(launch_tiles<JOIN_FUSED>(ctx, *plan));
    HMJ_CUDA(ctx, cudaMemcpyAsync(h, A.agg + 2, 8, cudaMemcpyDeviceToHost, ctx->stream));
    HMJ_CUDA(ctx, cudaMemcpyAsync(h + 1, A.collision, 4, cudaMemcpyDeviceToHost, ctx->stream));
    HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (!A.optimistic || (uint32_t)h[1] == 0) break;
    // two different keys under one hash: the optimistic counts were too high, join again exactly
  }
  A.optimistic = 0;
  *h_total = h[0];
  if (ctx->prof_on && ctx->prof_used >= 2 && h[0] <= cap)
    ctx->prof_bytes[ctx->prof_used / 2 - 1] += 8.0 * (double)h[0];
  return HMJ_OK;
}


// Samples a sorted hash column for equal-hash runs longer than RUN_PROBE rows: one thread per
// RUN_PROBE/2 rows compares h[i] with h[i+RUN_PROBE]; the hit count measures how many rows sit
// in such runs (every run of 1.5*RUN_PROBE rows or more is seen).
constexpr uint32_t RUN_PROBE = 4096;
__global__ void run_mass_kernel(const uint64_t* __restrict__ h, uint32_t n, unsigned long long* hits) {
  const uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * (RUN_PROBE / 2);
  const bool hit = i + RUN_PROBE < n && h[i] == h[i + RUN_PROBE];
  const unsigned b = __ballot_sync(0xffffffffu, hit);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(hits, (unsigned long long)__popc(b));
}

// The tile kernel walks, for every row of its first relation, that row's equal-hash run in the
// second one; a run is walked by one CTA, so the longest run of the second relation is the
// critical path.  Decide which relation goes second: the one with less mass in long runs.
int join_wants_swap(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s, bool* swap) {
  *swap = false;
  if (r->n <= RUN_PROBE && s->n <= RUN_PROBE) return HMJ_OK;
  uint64_t* scal = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_SCALARS, 8, &scal));
  unsigned long long* hits = reinterpret_cast<unsigned long long*>(scal + 4);
  HMJ_TRY(hmj_fill_async(ctx, hits, 0, 16, ctx->stream));
  const hmj_relation* rel[2] = {r, s};
  for (int k = 0; k < 2; ++k) {
    if (rel[k]->n <= RUN_PROBE) continue;
    const uint32_t samples = (uint32_t)hmj_div_up(rel[k]->n, RUN_PROBE / 2);
    run_mass_kernel<<<(samples + 255) / 256, 256, 0, ctx->stream>>>(rel[k]->d_hash, (uint32_t)rel[k]->n, hits + k);
    HMJ_LAUNCHED(ctx);
  }
  uint64_t* h = reinterpret_cast<uint64_t*>(ctx->pinned);
  HMJ_CUDA(ctx, cudaMemcpyAsync(h, hits, 16, cudaMemcpyDeviceToHost, ctx->stream));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *swap = h[1] > h[0];
  return HMJ_OK;
}

}  // namespace

// whole-operator helper (operator.cu): join two sorted relations into the context's result
// buffers, growing them when the first (optimistic) capacity was too small.  The pairs come out
// in unspecified order: when S holds the long equal-hash runs the relations trade places.
int hmj_impl_join_sorted(hmj_ctx* ctx, const hmj_relation* r_in, const hmj_relation* s_in,
                         uint64_t* h_num_pairs) {
  HMJ_TRY(check_relation(ctx, r_in));
  HMJ_TRY(check_relation(ctx, s_in));
  bool swap = false;
  HMJ_TRY(join_wants_swap(ctx, r_in, s_in, &swap));
  const hmj_relation* r = swap ? s_in : r_in;
  const hmj_relation* s = swap ? r_in : s_in;
  const int slot_r = swap ? WS_OUT_S : WS_OUT_R, slot_s = swap ? WS_OUT_R : WS_OUT_S;
  JoinPlan plan;
  HMJ_TRY(join_plan(ctx, r, s, &plan));
  // optimistic capacity: what the buffers already hold, at least max(|R|,|S|) pairs
  size_t cap = ctx->ws[WS_OUT_R].bytes / 4;
  const size_t guess = r->n > s->n ? r->n : s->n;
  if (cap < guess) cap = guess;
  uint32_t *out_r = nullptr, *out_s = nullptr;
  HMJ_TRY(hmj_ws(ctx, slot_r, cap, &out_r));
  HMJ_TRY(hmj_ws(ctx, slot_s, cap, &out_s));
  cap = ctx->ws[WS_OUT_R].bytes / 4;
  if (ctx->ws[WS_OUT_S].bytes / 4 < cap) cap = ctx->ws[WS_OUT_S].bytes / 4;
  uint64_t m = 0;
  HMJ_TRY(join_fused(ctx, &plan, out_r, out_s, cap, &m));
  if (m > cap) {
    HMJ_TRY(hmj_ws(ctx, slot_r, (size_t)m, &out_r));
    HMJ_TRY(hmj_ws(ctx, slot_s, (size_t)m, &out_s));
    uint64_t m2 = 0;
    HMJ_TRY(join_fused(ctx, &plan, out_r, out_s, m, &m2));
    m = m2;
  }
  ctx->last_pairs = m;
  *h_num_pairs = m;
  return HMJ_OK;
}

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
  HMJ_ARG(ctx, cap == 0 || (d_out_r && d_out_s));
  JoinPlan plan;
  HMJ_TRY(join_plan(ctx, r, s, &plan));
  HMJ_TRY(join_fused(ctx, &plan, d_out_r, d_out_s, cap, h_num_pairs));
  if (*h_num_pairs > cap)
    return hmj_fail(ctx, HMJ_ECAPACITY, "output capacity too small", __FILE__, __LINE__);
  return HMJ_OK;
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
  HMJ_TRY(hmj_fill_async(ctx, plan.args.agg, 0, 16, ctx->stream));
  HMJ_TRY(launch_tiles<JOIN_AGG>(ctx, plan));
  uint64_t* h = reinterpret_cast<uint64_t*>(ctx->pinned);
  HMJ_CUDA(ctx, cudaMemcpyAsync(h, plan.args.agg, 16, cudaMemcpyDeviceToHost, ctx->stream));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *h_num_pairs = h[0];
  *h_sum = h[1];
  return HMJ_OK;
}
