// radix_sort.cu -- K2 + K3: onesweep-style LSD radix sort of (u64 hash, u32 row-id) pairs.
//
// Stands in for the reference's MSD radix sort of (hash,key,value) tuples:
//   radix_non_inplace_par  radix_hash.h:351-423 (worker :294-346, recursion :122-292)
//   radix_inplace_seq/par  radix_hash.h:425-654
//   radix_int_inplace / radix_int_non_inplace  radix_sort.h:333-398,452-522
// The reference result that is pinned by its tests is "hash column fully ascending"; order
// inside an equal-hash run is unspecified there.  This sort is STABLE, so equal hashes come out
// in ascending input order, which makes the output canonical.
//
// Structure (all SoA, HBM resident):
//   K2  radix_histogram_kernel : one read of the key column -> 8 x 256 digit histograms
//   K2b radix_scan_kernel      : exclusive scan per digit -> global bin offsets, skip flags
//   K3  onesweep_pass_kernel   : per pass, one read + one write of the pairs.  Tiles are taken
//       from a dynamic counter; digits are ranked inside each warp with __match_any_sync
//       (warp-private shared-memory histograms), tile prefixes come from a decoupled look-back
//       chain over single-word (flag|epoch|count) descriptors; keys and row ids are staged in
//       shared memory in tile-sorted order so every digit bin is written as a contiguous run.
// Algorithmic bytes per pair: histogram 8; each executed pass 12 read + 12 written (the first
// pass of an iota-row-id sort reads 8).
#include <mutex>

#include "common.cuh"

namespace {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int NUM_PASSES = 64 / RADIX_BITS;

// ---------------------------------------------------------------- K2: histogram
constexpr int HIST_THREADS = 512;
constexpr int HIST_ITEMS = 8;  // keys per thread per iteration

__global__ void __launch_bounds__(HIST_THREADS)
radix_histogram_kernel(const uint64_t* __restrict__ keys, size_t n, uint32_t* __restrict__ g_hist) {
  __shared__ uint32_t s_hist[NUM_PASSES][RADIX];
  for (int i = threadIdx.x; i < NUM_PASSES * RADIX; i += HIST_THREADS) (&s_hist[0][0])[i] = 0;
  __syncthreads();
  const size_t chunk = (size_t)HIST_THREADS * HIST_ITEMS;
  for (size_t base = (size_t)blockIdx.x * chunk; base < n; base += (size_t)gridDim.x * chunk) {
    uint64_t k[HIST_ITEMS];
#pragma unroll
    for (int j = 0; j < HIST_ITEMS; ++j) {
      const size_t i = base + (size_t)j * HIST_THREADS + threadIdx.x;
      k[j] = i < n ? __ldcs(keys + i) : 0;
    }
#pragma unroll
    for (int j = 0; j < HIST_ITEMS; ++j) {
      const size_t i = base + (size_t)j * HIST_THREADS + threadIdx.x;
      if (i < n) {
#pragma unroll
        for (int p = 0; p < NUM_PASSES; ++p)
          atomicAdd(&s_hist[p][(k[j] >> (p * RADIX_BITS)) & (RADIX - 1)], 1u);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NUM_PASSES * RADIX; i += HIST_THREADS) {
    const uint32_t c = (&s_hist[0][0])[i];
    if (c) atomicAdd(g_hist + i, c);
  }
}

// One block per digit position: exclusive scan of its 256 bins (in place) plus the statistics
// the host planner reads back: g_stats[2p] = number of non-empty bins, g_stats[2p+1] = largest
// bin (== n means the digit is trivial: every key shares it and the pass can be skipped).
__global__ void __launch_bounds__(RADIX)
radix_scan_kernel(uint32_t* __restrict__ g_hist, uint32_t n32, uint32_t* __restrict__ g_stats) {
  __shared__ uint32_t s_warp[RADIX / 32];
  __shared__ uint32_t s_max[RADIX / 32];
  const int p = blockIdx.x, d = threadIdx.x, lane = d & 31, w = d >> 5;
  const uint32_t c = g_hist[p * RADIX + d];
  uint32_t x = c, mx = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if (lane == 31) s_warp[w] = x;
  if (lane == 0) s_max[w] = mx;
  const int nz = __syncthreads_count(c != 0);
  uint32_t pre = 0;
  for (int i = 0; i < w; ++i) pre += s_warp[i];
  g_hist[p * RADIX + d] = pre + x - c;
  if (d == 0) {
    uint32_t m = 0;
    for (int i = 0; i < RADIX / 32; ++i) m = max(m, s_max[i]);
    g_stats[2 * p] = (uint32_t)nz;
    g_stats[2 * p + 1] = m;
  }
}

// ---------------------------------------------------------------- K3: onesweep pass
// Look-back descriptor: one 64-bit word = status(2) | epoch(16) | count(46).  The epoch makes
// descriptors of earlier passes/sorts read as "not ready", so the chain needs no memset.
constexpr uint64_t LB_PARTIAL = 1ull << 62;
constexpr uint64_t LB_INCLUSIVE = 2ull << 62;
constexpr uint64_t LB_STATUS_MASK = 3ull << 62;
constexpr int LB_EPOCH_SHIFT = 46;
constexpr uint64_t LB_EPOCH_MASK = 0xFFFFull << LB_EPOCH_SHIFT;
constexpr uint64_t LB_VALUE_MASK = (1ull << LB_EPOCH_SHIFT) - 1;
constexpr int LB_WINDOW_DEFAULT = 4;

// Optional per-tile timeline of the pass (scripts/trace_pass.py; compiled only with -DHMJ_TRACE):
// 12 words per tile = globaltimer at {start, keys loaded, ranked, look-back start, look-back end,
// scattered, end}, look-back round trips, not-ready polls, SM id.
#ifdef HMJ_TRACE
__device__ unsigned long long* g_trace = nullptr;
__device__ unsigned long long g_trace_tiles = 0;
__device__ __forceinline__ unsigned long long trace_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TRACE(slot) do { if (tid == 0 && g_trace && tile < g_trace_tiles) g_trace[(size_t)tile * 12 + (slot)] = trace_now(); } while (0)
#define TRACE_VAL(slot, v) do { if (tid == 0 && g_trace && tile < g_trace_tiles) g_trace[(size_t)tile * 12 + (slot)] = (v); } while (0)
#else
#define TRACE(slot) do {} while (0)
#define TRACE_VAL(slot, v) do {} while (0)
#endif

__device__ __forceinline__ void lb_store(uint64_t* p, uint64_t v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t lb_load(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

template <int THREADS, int ITEMS, int MIN_BLOCKS = 1>
struct OnesweepCfg {
  static constexpr int WARPS = THREADS / 32;
  static constexpr int TILE = THREADS * ITEMS;
  static constexpr size_t SMEM_BYTES =
      (size_t)TILE * 8 + (size_t)TILE * 4 + (size_t)WARPS * RADIX * 4 + 3 * RADIX * 4 + 64;
};

// keys_in/vals_in -> keys_out/vals_out, stable on digit (key >> shift) & 255.
// vals_in == nullptr means the payload is the element's global index (row id iota).
template <int THREADS, int ITEMS, int MIN_BLOCKS, int RANK_MODE, int LB_WINDOW>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
onesweep_pass_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                     uint32_t n, int shift, const uint32_t* __restrict__ bin_offsets,
                     uint64_t* __restrict__ lookback, uint32_t* __restrict__ tile_counter,
                     uint32_t epoch, uint32_t iota_base) {
  using Cfg = OnesweepCfg<THREADS, ITEMS>;
  constexpr int WARPS = Cfg::WARPS;
  constexpr int TILE = Cfg::TILE;
  static_assert(THREADS >= RADIX, "one thread per digit is needed for the look-back");

  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem_raw);
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys + TILE);
  uint32_t* s_whist = s_vals + TILE;             // [WARPS][RADIX] warp-private counts/offsets
  uint32_t* s_tile_off = s_whist + WARPS * RADIX;  // [RADIX] start of digit inside the tile
  uint32_t* s_gbase = s_tile_off + RADIX;          // [RADIX] global start of digit minus tile_off
  uint32_t* s_scan = s_gbase + RADIX;              // [RADIX/32] + tile id
  uint32_t* s_tile_id = s_scan + RADIX / 32 + 1;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) *s_tile_id = atomicAdd(tile_counter, 1u);
  for (int i = tid; i < WARPS * RADIX; i += THREADS) s_whist[i] = 0;
  __syncthreads();
  const uint32_t tile = *s_tile_id;
  const uint32_t tile_base = tile * (uint32_t)TILE;
  const uint32_t valid = min((uint32_t)TILE, n - tile_base);
  TRACE(0);

  // ---- load keys, warp-striped: item k of lane l sits at warp_base + k*32 + l
  const uint32_t warp_base = tile_base + warp * (32 * ITEMS);
  uint64_t key[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const uint32_t idx = warp_base + k * 32 + lane;
    key[k] = (idx < n) ? __ldcs(keys_in + idx) : ~0ull;
  }

#ifdef HMJ_TRACE
  if (key[ITEMS - 1] == 0x123456789abcdefull) return;  // forces the loads to complete here
  TRACE(1);
#endif
  // ---- rank digits inside the warp (stable in item order k, lane)
  uint32_t* whist = s_whist + warp * RADIX;
  uint16_t rank[ITEMS];
  const uint32_t lt_mask = (1u << lane) - 1;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const uint32_t d = (uint32_t)(key[k] >> shift) & (RADIX - 1);
    uint32_t peers;
    // RANK_MODE = how many of every 4 items use MATCH.ANY (ADU pipe, ~65 cycles per warp
    // instruction per SM, profiles/r1a_*); the rest use one ballot per digit bit (ALU pipe).
    // Mixing keeps both pipes busy.  4 = all MATCH (round-1a kernel), 0 = all ballots.
    if ((k & 3) < RANK_MODE) {
      peers = __match_any_sync(0xffffffffu, d);
    } else {
      peers = 0xffffffffu;
#pragma unroll
      for (int b = 0; b < RADIX_BITS; ++b) {
        asm("{\n\t"
            ".reg .pred p;\n\t"
            ".reg .b32 m;\n\t"
            "and.b32 m, %1, %2;\n\t"
            "setp.ne.u32 p, m, 0;\n\t"
            "vote.sync.ballot.b32 m, p, 0xffffffff;\n\t"
            "@!p not.b32 m, m;\n\t"
            "and.b32 %0, %0, m;\n\t"
            "}"
            : "+r"(peers)
            : "r"(d), "r"(1u << b));
      }
    }
    // the lowest peer adds the group's count to the warp-private bin; shared-memory atomics of
    // one warp are performed in issue order, so item k+1 sees item k's count without a
    // __syncwarp and the 16 chains can overlap
    const int leader = __ffs(peers) - 1;
    uint32_t old = 0;
    if (lane == leader) old = atomicAdd(&whist[d], (uint32_t)__popc(peers));
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[k] = (uint16_t)(old + __popc(peers & lt_mask));
  }
  __syncwarp();
  __syncthreads();

  TRACE(2);
  // ---- per digit: exclusive offsets over warps, tile totals (published as PARTIAL at once),
  // scan over digits
  uint32_t pub = 0, tile_off = 0;
  const uint64_t tag = (uint64_t)epoch << LB_EPOCH_SHIFT;
  uint64_t* my_desc = lookback + (size_t)tile * RADIX + (tid & (RADIX - 1));
  if (tid < RADIX) {
    uint32_t sum = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = s_whist[w * RADIX + tid];
      s_whist[w * RADIX + tid] = sum;
      sum += c;
    }
    const uint32_t count = sum;
    // out-of-range items were given digit 255 and rank last inside it; they are not data.
    pub = count;
    if (tid == RADIX - 1) pub -= (uint32_t)TILE - valid;
    if (tile != 0) lb_store(my_desc, LB_PARTIAL | tag | pub);
    // inclusive scan of tile counts across the 256 digits
    uint32_t x = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_scan[warp] = x;
    // (only warps 0..7 are in this branch; they synchronise on a named barrier)
    asm volatile("bar.sync 1, %0;" ::"n"(RADIX));
    uint32_t pre = 0;
    for (int i = 0; i < warp; ++i) pre += s_scan[i];
    tile_off = pre + x - count;
    s_tile_off[tid] = tile_off;
  }
  __syncthreads();

  // ---- decoupled look-back over earlier tiles of this digit, LB_WINDOW descriptors per round
  // trip.  Tiles start every ~30 ns while one L2 round trip is several hundred ns, so a serial
  // walk (one descriptor per trip) cannot keep up once a tile is delayed: the window must cover
  // the tiles that start during one trip (profiles/r1c: 30 % of the stall samples sat here).
  TRACE(3);
  if (tid < RADIX) {
    uint32_t excl = 0;
    unsigned long long trips = 0, polls = 0;
    if (tile != 0) {
      int64_t prev = (int64_t)tile - 1;
      bool done = false;
      while (!done) {
        ++trips;
        uint64_t v[LB_WINDOW];
#pragma unroll
        for (int j = 0; j < LB_WINDOW; ++j)
          v[j] = (prev - j >= 0) ? lb_load(lookback + (size_t)(prev - j) * RADIX + tid) : 0;
        int used = 0;
#pragma unroll
        for (int j = 0; j < LB_WINDOW; ++j) {
          if (done || used != j) continue;
          if ((v[j] & LB_EPOCH_MASK) != tag || (v[j] & LB_STATUS_MASK) == 0) continue;  // not ready
          excl += (uint32_t)(v[j] & LB_VALUE_MASK);
          ++used;
          if ((v[j] & LB_STATUS_MASK) == LB_INCLUSIVE) done = true;
        }
        prev -= used;
        if (used == 0) ++polls;
      }
    }
    TRACE_VAL(7, trips);
    TRACE_VAL(8, polls);
    TRACE(4);
    lb_store(my_desc, LB_INCLUSIVE | tag | (uint64_t)(excl + pub));
    s_gbase[tid] = bin_offsets[tid] + excl - tile_off;
  }

  // ---- scatter keys into tile-sorted order in shared memory (tile-local offsets only)
  {
    uint16_t pos[ITEMS];
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const uint32_t d = (uint32_t)(key[k] >> shift) & (RADIX - 1);
      const uint32_t p = s_tile_off[d] + whist[d] + rank[k];
      pos[k] = (uint16_t)p;
      s_keys[p] = key[k];
    }
    // payloads (loaded late to keep registers low)
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const uint32_t idx = warp_base + k * 32 + lane;
      uint32_t v = idx + iota_base;
      if (vals_in != nullptr && idx < n) v = __ldcs(vals_in + idx);
      s_vals[pos[k]] = v;
    }
  }
  __syncthreads();

  TRACE(5);
  // ---- write out: consecutive threads write consecutive slots of a digit's run
#pragma unroll 4
  for (uint32_t i = tid; i < valid; i += THREADS) {
    const uint64_t kk = s_keys[i];
    const uint32_t d = (uint32_t)(kk >> shift) & (RADIX - 1);
    const uint32_t dst = s_gbase[d] + i;
    keys_out[dst] = kk;
    vals_out[dst] = s_vals[i];
  }
#ifdef HMJ_TRACE
  TRACE(6);
  unsigned smid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  TRACE_VAL(9, (unsigned long long)smid);
#endif
}

}  // namespace

#ifdef HMJ_TRACE
extern "C" int hmj_debug_trace(unsigned long long* d_buf, unsigned long long tiles) {
  cudaError_t e = cudaMemcpyToSymbol(g_trace, &d_buf, sizeof d_buf);
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_trace_tiles, &tiles, sizeof tiles);
  return e == cudaSuccess ? 0 : -2;
}
#endif

// ---------------------------------------------------------------- host driver
namespace {

__global__ void iota_kernel(uint32_t* __restrict__ out, uint32_t n) {
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = i;
}

template <int THREADS, int ITEMS, int MIN_BLOCKS, int RANK_MODE = 0, int LBW = LB_WINDOW_DEFAULT>
int launch_pass(hmj_ctx* ctx, const uint64_t* kin, const uint32_t* vin, uint64_t* kout,
                uint32_t* vout, uint32_t n, int shift, const uint32_t* bins, uint64_t* lookback,
                uint32_t* counter, uint32_t epoch, uint32_t iota_base = 0) {
  using Cfg = OnesweepCfg<THREADS, ITEMS>;
  // per-device attribute; cheap enough to set on every launch
  HMJ_CUDA(ctx, cudaFuncSetAttribute(onesweep_pass_kernel<THREADS, ITEMS, MIN_BLOCKS, RANK_MODE, LBW>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)Cfg::SMEM_BYTES));
  const uint32_t tiles = (uint32_t)hmj_div_up(n, Cfg::TILE);
  cudaEvent_t ev_after;
  HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_PASS, (double)n * (vin ? 24.0 : 20.0), &ev_after));
  onesweep_pass_kernel<THREADS, ITEMS, MIN_BLOCKS, RANK_MODE, LBW><<<tiles, THREADS, Cfg::SMEM_BYTES, ctx->stream>>>(
      kin, vin, kout, vout, n, shift, bins, lookback, counter, epoch, iota_base);
  HMJ_LAUNCHED(ctx);
  hmj_prof_end(ctx, ev_after);
  return HMJ_OK;
}

int sort_cfg_from_env() {
  // tuning knob for scripts/sort_sweep.py; read per call so one process can sweep
  const char* e = getenv("HMJ_SORT_CFG");
  return e ? atoi(e) : 0;
}

int tile_items_for_cfg(int cfg) {
  switch (cfg) {
    case 3: case 8: case 9: return OnesweepCfg<384, 16>::TILE;
    case 4: case 6: case 7: return OnesweepCfg<256, 24>::TILE;
    default: return OnesweepCfg<256, 16>::TILE;
  }
}

}  // namespace

// ---------------------------------------------------------------- bucket finish (hybrid sort)
// After the scatter passes over the top digits the array is ordered on the prefix
// key >> shift, stable inside a bucket.  This kernel finishes every bucket by rank-by-counting
// on the full key (ties keep their order, so the whole sort stays stable): it stands in for the
// reference's small-bucket insertion sort (radix_hash.h:86-120,147-150).  A CTA owns the
// buckets that START in its tile and sees FIN_HALO elements past the tile end; a bucket that
// does not fit in that window or exceeds FIN_MAX_BUCKET raises *flag and the host re-sorts
// with scatter passes over every digit.  One read + one write of the pairs.
namespace {
constexpr int FIN_THREADS = 256;
constexpr int FIN_TILE = 2048;
constexpr int FIN_HALO = 512;
constexpr int FIN_MAX_BUCKET = 512;
constexpr int FIN_SPAN = FIN_TILE + FIN_HALO;

// IN PLACE: only the key column is read (8 B per pair); an element moves only when its rank
// inside the bucket differs from its position, and then its row id moves with it.  With the
// planner's prefix (buckets of ~1 element on average) almost nothing moves.
//
// fast path data: one 32-bit word per element = (31 key bits just below the prefix) << 1 | head
// flag (first element of its bucket).  Ranking walks compare these words; two bucket-mates that
// agree on all 31 bits but differ in the full key (possible only when shift > 31) send the
// tile to the slow path, which repeats the walks on the full 64-bit keys.
__global__ void __launch_bounds__(FIN_THREADS)
bucket_finish_kernel(uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t n,
                     int shift, uint32_t* __restrict__ flag) {
  constexpr int ITERS = (FIN_SPAN + FIN_THREADS - 1) / FIN_THREADS;
  constexpr uint32_t NONE = 0xffffffffu;
  __shared__ uint64_t s_k[FIN_SPAN + 1];   // [0] = element before the tile
  __shared__ uint32_t s_sub[FIN_SPAN + 1]; // [avail] = sentinel head
  __shared__ uint16_t s_src[FIN_SPAN + 1]; // source index of the element ranked at a position
  __shared__ uint32_t s_tie;
  const uint32_t tile_start = blockIdx.x * (uint32_t)FIN_TILE;
  const uint32_t avail = min((uint32_t)FIN_SPAN, n - tile_start);
  const int tid = threadIdx.x;
  const bool has_pred = tile_start != 0;
  if (tid == 0) {
    s_k[0] = has_pred ? keys[tile_start - 1] : 0;
    s_tie = 0;
  }
  for (uint32_t i = tid; i < avail; i += FIN_THREADS) s_k[1 + i] = keys[tile_start + i];
  __syncthreads();
  const int sub_shift = shift > 31 ? shift - 31 : 0;
  const uint64_t sub_mask = shift >= 31 ? 0x7fffffffull : ((1ull << shift) - 1);
  for (uint32_t i = tid; i <= avail; i += FIN_THREADS) {
    uint32_t w = 1u;  // sentinel
    if (i < avail) {
      const uint64_t key = s_k[1 + i], prev = s_k[i];
      const bool head = (i == 0) || (shift < 64 && ((key ^ prev) >> shift) != 0);
      w = ((uint32_t)((key >> sub_shift) & sub_mask) << 1) | (head ? 1u : 0u);
    }
    s_sub[i] = w;
    s_src[i] = 0xffff;
  }
  __syncthreads();
  // the bucket at the tile start belongs to the previous tile when it began there
  const bool straddle = has_pred && (shift >= 64 || ((s_k[1] ^ s_k[0]) >> shift) == 0);
  const bool more_after = tile_start + avail < n;

  uint32_t my_pos[ITERS];  // new position (tile-relative) of an element that has to move
#pragma unroll
  for (int it = 0; it < ITERS; ++it) {
    const uint32_t i = it * FIN_THREADS + tid;
    my_pos[it] = NONE;
    if (i >= avail) continue;
    const uint32_t me = s_sub[i];
    if ((me & 1u) && (s_sub[i + 1] & 1u)) continue;  // bucket of one element
    const uint32_t me_le = me | 1u, me_lt = me & ~1u;
    uint32_t rank = 0, o = me, j = i;
    while (!(o & 1u)) {
      --j;
      o = s_sub[j];
      rank += (o <= me_le) ? 1u : 0u;
    }
    const uint32_t start = j;
    if (start >= (uint32_t)FIN_TILE || (start == 0 && straddle)) continue;  // not owned here
    uint32_t e = i + 1;
    o = s_sub[e];
    while (!(o & 1u)) {
      rank += (o < me_lt) ? 1u : 0u;
      ++e;
      o = s_sub[e];
    }
    if ((e == avail && more_after) || e - start > (uint32_t)FIN_MAX_BUCKET) {
      if (i == start) atomicOr(flag, 1u);
      continue;
    }
    my_pos[it] = start + rank;
    if (shift > 31) s_src[start + rank] = (uint16_t)i;
  }
  // tie check: my successor in ranked order, if a bucket-mate with the same 31 bits, must
  // also have the same full key
  if (shift > 31) {
    __syncthreads();
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      const uint32_t i = it * FIN_THREADS + tid;
      const uint32_t p = my_pos[it];
      if (p == NONE) continue;
      const uint32_t nxt = s_src[p + 1];
      if (nxt == 0xffff) continue;
      if (((s_sub[nxt] ^ s_sub[i]) >> 1) == 0 &&
          (shift >= 64 || ((s_k[1 + nxt] ^ s_k[1 + i]) >> shift) == 0) &&
          s_k[1 + nxt] != s_k[1 + i])
        s_tie = 1;
    }
    __syncthreads();
    if (s_tie != 0) {
      // slow path: same ranking on the full keys
      const uint64_t low_mask = shift >= 64 ? ~0ull : ((1ull << shift) - 1);
#pragma unroll
      for (int it = 0; it < ITERS; ++it) {
        const uint32_t i = it * FIN_THREADS + tid;
        my_pos[it] = NONE;
        if (i >= avail) continue;
        const uint64_t key = s_k[1 + i];
        const uint64_t lo = key & ~low_mask, hi = key | low_mask;  // key range of my bucket
        uint32_t rank = 0;
        int j = (int)i - 1;
        while (j >= 0) {
          const uint64_t o = s_k[1 + j];
          if (o < lo) break;
          rank += (o <= key) ? 1u : 0u;
          --j;
        }
        const uint32_t start = (uint32_t)(j + 1);
        if (start >= (uint32_t)FIN_TILE) continue;
        if (j < 0 && has_pred && s_k[0] >= lo) continue;
        uint32_t e = i + 1;
        while (e < avail) {
          const uint64_t o = s_k[1 + e];
          if (o > hi) break;
          rank += (o < key) ? 1u : 0u;
          ++e;
        }
        if ((e == avail && more_after) || e - start > (uint32_t)FIN_MAX_BUCKET) continue;  // flagged above
        my_pos[it] = start + rank;
      }
    }
  }
  // ---- move phase: row ids of every moving element are read before anything is written
  // (buckets are owned by exactly one CTA, so nobody else writes these positions)
  uint32_t my_val[ITERS];
  bool moving = false;
#pragma unroll
  for (int it = 0; it < ITERS; ++it) {
    const uint32_t i = it * FIN_THREADS + tid;
    if (my_pos[it] == i) my_pos[it] = NONE;  // already in place
    if (my_pos[it] != NONE) {
      my_val[it] = vals[tile_start + i];
      moving = true;
    }
  }
  if (!__syncthreads_or(moving)) return;
#pragma unroll
  for (int it = 0; it < ITERS; ++it) {
    const uint32_t i = it * FIN_THREADS + tid;
    if (my_pos[it] == NONE) continue;
    keys[tile_start + my_pos[it]] = s_k[1 + i];
    vals[tile_start + my_pos[it]] = my_val[it];
  }
}

struct DigitStats { uint32_t nz, maxbin; };

int hybrid_from_env() {
  const char* e = getenv("HMJ_SORT_HYBRID");
  return e ? atoi(e) : 2;
}

}  // namespace

// Plans and runs one sort.  Steps alternate between (keys_out,vals_out) and the workspace
// ping-pong pair; the last step always lands in the output pair.
int hmj_impl_sort(hmj_ctx* ctx, const uint64_t* keys_in, const uint32_t* vals_in, size_t n,
                  uint64_t* keys_out, uint32_t* vals_out) {
  HMJ_ARG(ctx, n < (size_t(1) << 32) - (size_t(1) << 16));
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, keys_in && keys_out && vals_out);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const uint32_t n32 = (uint32_t)n;
  const int cfg = sort_cfg_from_env();
  const int tile_items = tile_items_for_cfg(cfg);
  const size_t tiles = hmj_div_up(n, tile_items);

  const int hyb = hybrid_from_env();
  const uint32_t seg_cap = (uint32_t)hmj_segment_sort_cap();
  // hashes with implicit row ids: try the uniform-hash path first (no histogram, no stable
  // passes); it declines or fails on skewed / heavily duplicated keys and leaves the input intact
  // (a join whose sample already found this relation skewed says so: no attempt to throw away)
  const bool known_skewed = ctx->next_sort_skewed;
  ctx->next_sort_skewed = false;
  if (vals_in == nullptr && hyb == 2 && !known_skewed) {
    int ok = 0;
    HMJ_TRY(hmj_sort_uniform_try(ctx, keys_in, n, 0, keys_out, vals_out, &ok));
    if (ok) {
      ctx->tm.sort_uniform += 1;
      return HMJ_OK;
    }
  }

  // scratch: [8*256 bins][16 stats][24 tile counters][8 flags]
  constexpr int MAX_LAUNCHED_PASSES = 24;
  uint32_t* d_hist = nullptr;
  const size_t hist_words = NUM_PASSES * RADIX + 2 * NUM_PASSES + MAX_LAUNCHED_PASSES + 8;
  HMJ_TRY(hmj_ws(ctx, WS_HIST, hist_words, &d_hist));
  uint32_t* d_stats = d_hist + NUM_PASSES * RADIX;
  uint32_t* d_counters = d_stats + 2 * NUM_PASSES;
  uint32_t* d_flag = d_counters + MAX_LAUNCHED_PASSES;
  HMJ_TRY(hmj_fill_async(ctx, d_hist, 0, hist_words * sizeof(uint32_t), st));

  const size_t chunk = (size_t)HIST_THREADS * HIST_ITEMS;
  const int hgrid = (int)min(hmj_div_up(n, chunk), (size_t)ctx->num_sms * 4);
  cudaEvent_t ev_hist;
  HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_HIST, (double)n * 8.0, &ev_hist));
  radix_histogram_kernel<<<hgrid, HIST_THREADS, 0, st>>>(keys_in, n, d_hist);
  HMJ_LAUNCHED(ctx);
  hmj_prof_end(ctx, ev_hist);
  radix_scan_kernel<<<NUM_PASSES, RADIX, 0, st>>>(d_hist, n32, d_stats);
  HMJ_LAUNCHED(ctx);

  // ---- plan (host): which digits need a scatter pass, and whether a prefix sort + bucket
  // finish is enough.  Small sorts skip the read-back and run every pass.
  DigitStats stats[NUM_PASSES];
  for (int p = 0; p < NUM_PASSES; ++p) stats[p] = {RADIX, 0};
  const bool planned = hyb == 2 ? n > seg_cap : n >= (size_t(1) << 15);
  if (planned) {
    uint32_t* h = reinterpret_cast<uint32_t*>(ctx->pinned);
    HMJ_CUDA(ctx, cudaMemcpyAsync(h, d_stats, 2 * NUM_PASSES * sizeof(uint32_t),
                                  cudaMemcpyDeviceToHost, st));
    HMJ_CUDA(ctx, cudaStreamSynchronize(st));
    for (int p = 0; p < NUM_PASSES; ++p) stats[p] = {h[2 * p], h[2 * p + 1]};
  }
  bool trivial[NUM_PASSES];
  for (int p = 0; p < NUM_PASSES; ++p) trivial[p] = planned && stats[p].maxbin == n32;

  int digits[NUM_PASSES], nd = 0;  // scatter passes of the first attempt, ascending
  int finish_shift = -1;           // >= 0: finish on the buckets of key >> finish_shift
  int finish_kind = 0;             // 1 = bucket finish (in place, small buckets), 2 = segment sort
  int seg_group_shift = -1;        // segment sort: >= 0 = windows of 2^shift prefix values
  const uint32_t seg_stride = seg_cap - 1024;  // mode B: segments up to 1024 pairs fit
  if (hyb == 2 && n <= seg_cap) {
    finish_kind = 2;  // the whole array is one window
    finish_shift = 64;
  } else if (hyb == 2 && planned) {
    // Scatter on the top digits until the prefix cuts the array into segments the segment sort
    // takes in one window: either one window per prefix value (dense digits, mean + 8 sigma
    // within the window capacity) or several small segments per window.
    double max_avg_a = seg_cap;
    while (max_avg_a + 8.0 * sqrt(max_avg_a) > (double)seg_cap) max_avg_a -= 16.0;
    const double max_avg_b = 640.0;
    double eff = 0;
    int q = 0;
    bool skewed = false, dense = true;
    while (q < NUM_PASSES) {
      const double avg = (double)n / exp2(eff);
      if (dense && q > 0 && avg <= max_avg_a && avg > max_avg_b) {
        // windows of g prefix values, g the largest power of two with g * avg <= max_avg_a
        seg_group_shift = (int)floor(log2(max_avg_a / avg));
        seg_group_shift = min(seg_group_shift, RADIX_BITS * q);
        if (RADIX_BITS * q - seg_group_shift <= 20) break;
        seg_group_shift = -1;
      }
      if (avg <= max_avg_b) break;
      const int d = NUM_PASSES - 1 - q;
      if (!trivial[d]) {
        const double mean = (double)n / stats[d].nz;
        if ((double)stats[d].maxbin > mean + 8.0 * sqrt(mean) + 64.0) skewed = true;
        eff += log2((double)stats[d].nz);
        if (stats[d].nz != RADIX) dense = false;
      } else {
        dense = false;
      }
      ++q;
    }
    if (!skewed && q < NUM_PASSES) {
      finish_kind = 2;
      finish_shift = 64 - RADIX_BITS * q;
      for (int d = NUM_PASSES - q; d < NUM_PASSES; ++d)
        if (!trivial[d]) digits[nd++] = d;
    }
  } else if (hyb == 1 && (planned || n <= (size_t)FIN_MAX_BUCKET)) {
    // include top digits until the prefix separates the keys into buckets of <= ~16..32
    const double need = n <= (size_t)FIN_MAX_BUCKET / 2 ? 0.0 : floor(log2((double)n)) - 3.0;
    double eff = 0;
    int q = 0, nontrivial = 0;
    bool skewed = false;
    while (q < NUM_PASSES && eff < need) {
      const int d = NUM_PASSES - 1 - q;
      if (!trivial[d]) {
        const double mean = (double)n / stats[d].nz;
        if ((double)stats[d].maxbin > mean + 8.0 * sqrt(mean) + 64.0) skewed = true;
        eff += log2((double)stats[d].nz);
        ++nontrivial;
      }
      ++q;
    }
    if (!(skewed || eff < need || nontrivial >= NUM_PASSES - 1)) {
      finish_shift = RADIX_BITS * (NUM_PASSES - q);
      if (finish_shift != 0) finish_kind = 1;  // else the prefix is the whole key: plain LSD
      else finish_shift = -1;
      for (int d = NUM_PASSES - q; d < NUM_PASSES; ++d)
        if (!trivial[d]) digits[nd++] = d;
    }
  }
  if (finish_kind == 0 && finish_shift < 0 && nd == 0) {
    for (int d = 0; d < NUM_PASSES; ++d)
      if (!trivial[d]) digits[nd++] = d;
  }
  if (finish_kind == 0) finish_shift = -1;

  struct Buf { uint64_t* k; uint32_t* v; };
  Buf bufs[2] = {{keys_out, vals_out}, {nullptr, nullptr}};
  const bool aliased = (keys_in == keys_out) || (vals_in != nullptr && vals_in == vals_out);
  int steps = nd;  // the bucket finish works in place
  if (steps == 0 && finish_shift < 0) {  // nothing to do: copy through
    if (keys_in != keys_out)
      HMJ_CUDA(ctx, cudaMemcpyAsync(keys_out, keys_in, n * 8, cudaMemcpyDeviceToDevice, st));
    if (vals_in == nullptr) {
      iota_kernel<<<(int)min(hmj_div_up(n, 256), (size_t)ctx->num_sms * 8), 256, 0, st>>>(vals_out, n32);
      HMJ_LAUNCHED(ctx);
    } else if (vals_in != vals_out) {
      HMJ_CUDA(ctx, cudaMemcpyAsync(vals_out, vals_in, n * 4, cudaMemcpyDeviceToDevice, st));
    }
    return HMJ_OK;
  }
  HMJ_TRY(hmj_ws(ctx, WS_SORT_KEYS_A, n, &bufs[1].k));
  HMJ_TRY(hmj_ws(ctx, WS_SORT_VALS_A, n, &bufs[1].v));

  uint64_t* d_lb = nullptr;
  if (nd > 0 || finish_shift >= 0) {
    HMJ_TRY(hmj_ws(ctx, WS_LOOKBACK, tiles * RADIX, &d_lb));
    if (ctx->lb_seen_ptr != d_lb || ctx->lb_seen_bytes != ctx->ws[WS_LOOKBACK].bytes ||
        ctx->lb_epoch + 3 * NUM_PASSES >= 0xFFFF) {
      HMJ_TRY(hmj_fill_async(ctx, d_lb, 0, ctx->ws[WS_LOOKBACK].bytes, st));
      ctx->lb_seen_ptr = d_lb;
      ctx->lb_seen_bytes = ctx->ws[WS_LOOKBACK].bytes;
      ctx->lb_epoch = 0;
    }
  }

  int counter_idx = 0;
  // runs scatter passes over digits[0..cnt) starting from (src_k,src_v); step j writes
  // bufs[first_dst ^ (j & 1)].  Returns the index of the buffer holding the result.
  auto run_passes = [&](const uint64_t* src_k, const uint32_t* src_v, const int* dg, int cnt,
                        int first_dst, int* where) -> int {
    const uint64_t* cur_k = src_k;
    const uint32_t* cur_v = src_v;
    int w = first_dst ^ 1;
    for (int j = 0; j < cnt; ++j) {
      w = first_dst ^ (j & 1);
      uint64_t* dk = bufs[w].k;
      uint32_t* dv = bufs[w].v;
      const int p = dg[j];
      const uint32_t epoch = ++ctx->lb_epoch;
      const uint32_t* bins = d_hist + p * RADIX;
      uint32_t* counter = d_counters + (counter_idx++);
      int rc;
#define HMJ_PASS(T, I, B, R, ...) \
  rc = launch_pass<T, I, B, R, ##__VA_ARGS__>(ctx, cur_k, cur_v, dk, dv, n32, p * RADIX_BITS, bins, d_lb, counter, epoch)
      switch (cfg) {
        case 1: HMJ_PASS(256, 16, 3, 1); break;
        case 2: HMJ_PASS(256, 16, 3, 2); break;
        case 3: HMJ_PASS(384, 16, 2, 0); break;
        case 4: HMJ_PASS(256, 24, 2, 0); break;
        case 5: HMJ_PASS(256, 16, 3, 4); break;
        case 6: HMJ_PASS(256, 24, 2, 1); break;
        case 7: HMJ_PASS(256, 24, 2, 2); break;
        case 8: HMJ_PASS(384, 16, 2, 1); break;
        case 9: HMJ_PASS(384, 16, 2, 2); break;
        case 10: HMJ_PASS(256, 16, 3, 0, 1); break;
        case 11: HMJ_PASS(256, 16, 3, 0, 2); break;
        case 13: HMJ_PASS(256, 16, 3, 0, 8); break;
        case 14: HMJ_PASS(256, 16, 3, 0, 16); break;
        default: HMJ_PASS(256, 16, 3, 0); break;
      }
#undef HMJ_PASS
      HMJ_TRY(rc);
      cur_k = dk;
      cur_v = dv;
    }
    *where = w;
    return HMJ_OK;
  };

  // first destination chosen so that the last step lands in bufs[0]; an aliased input cannot
  // be overwritten by the step that reads it, so it goes to the workspace first
  int first_dst = (steps - 1) & 1;
  if (aliased && first_dst == 0) first_dst = 1;
  int where = 0;
  if (nd > 0) {
    HMJ_TRY(run_passes(keys_in, vals_in, digits, nd, first_dst, &where));
    ctx->tm.sort_passes += (uint32_t)nd;
  } else {
    // no scatter pass (tiny input, or the prefix digits are all trivial): the pairs go to the
    // output columns as they are and the finish orders them there
    if (keys_in != keys_out)
      HMJ_CUDA(ctx, cudaMemcpyAsync(keys_out, keys_in, n * 8, cudaMemcpyDeviceToDevice, st));
    if (vals_in == nullptr) {
      iota_kernel<<<(int)min(hmj_div_up(n, 256), (size_t)ctx->num_sms * 8), 256, 0, st>>>(vals_out, n32);
      HMJ_LAUNCHED(ctx);
    } else if (vals_in != vals_out) {
      HMJ_CUDA(ctx, cudaMemcpyAsync(vals_out, vals_in, n * 4, cudaMemcpyDeviceToDevice, st));
    }
    where = 0;
  }

  if (finish_shift >= 0) {
    cudaEvent_t ev_fin;
    HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_FINISH, (double)n * (finish_kind == 2 ? 24.0 : 8.0), &ev_fin));
    if (finish_kind == 2) {
      HMJ_TRY(hmj_launch_segment_sort(ctx, bufs[where].k, bufs[where].v, n32, 64 - finish_shift,
                                      seg_group_shift, seg_stride, d_flag));
    } else {
      bucket_finish_kernel<<<(unsigned)hmj_div_up(n, FIN_TILE), FIN_THREADS, 0, st>>>(
          bufs[where].k, bufs[where].v, n32, finish_shift, d_flag);
      HMJ_LAUNCHED(ctx);
    }
    hmj_prof_end(ctx, ev_fin);
    ctx->tm.sort_finishes += 1;
    uint32_t* h = reinterpret_cast<uint32_t*>(ctx->pinned);
    HMJ_CUDA(ctx, cudaMemcpyAsync(h, d_flag, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    HMJ_CUDA(ctx, cudaStreamSynchronize(st));
    if (h[0] != 0) {
      // A bucket was too large.  What the finish left behind is still a permutation of the
      // input that keeps equal keys in input order, so stable scatter passes over every
      // non-trivial digit, started from there, give the same result as a sort of the input.
      ctx->tm.sort_fallbacks += 1;
      int all[NUM_PASSES], na = 0;
      for (int d = 0; d < NUM_PASSES; ++d)
        if (!trivial[d]) all[na++] = d;
      if (na > 0) {
        const int src = where;
        HMJ_TRY(run_passes(bufs[src].k, bufs[src].v, all, na, src ^ 1, &where));
        ctx->tm.sort_passes += (uint32_t)na;
      }
      // na == 0: every key is equal, the order is already right
    }
  }
  if (where != 0) {
    HMJ_CUDA(ctx, cudaMemcpyAsync(keys_out, bufs[1].k, n * 8, cudaMemcpyDeviceToDevice, st));
    HMJ_CUDA(ctx, cudaMemcpyAsync(vals_out, bufs[1].v, n * 4, cudaMemcpyDeviceToDevice, st));
  }
  return HMJ_OK;
}

extern "C" int hmj_sort_pairs(hmj_ctx* ctx, const uint64_t* d_hash, const uint32_t* d_rowid,
                              size_t n, uint64_t* d_hash_out, uint32_t* d_rowid_out) {
  if (!ctx) return HMJ_EARG;
  return hmj_impl_sort(ctx, d_hash, d_rowid, n, d_hash_out, d_rowid_out);
}

extern "C" int hmj_sort_int_pairs(hmj_ctx* ctx, const uint64_t* d_keys, const uint32_t* d_vals,
                                  size_t n, uint64_t* d_keys_out, uint32_t* d_vals_out) {
  if (!ctx) return HMJ_EARG;
  return hmj_impl_sort(ctx, d_keys, d_vals, n, d_keys_out, d_vals_out);
}

extern "C" int hmj_hash_sort_u64(hmj_ctx* ctx, const uint64_t* d_keys, size_t n,
                                 uint64_t* d_hash_out, uint32_t* d_rowid_out) {
  if (!ctx) return HMJ_EARG;
  // std::hash<integral> is the identity: the key column IS the hash column, and the first
  // scatter pass generates the row ids, so no separate hash kernel runs.
  return hmj_impl_sort(ctx, d_keys, nullptr, n, d_hash_out, d_rowid_out);
}

extern "C" int hmj_hash_sort_strings(hmj_ctx* ctx, const uint64_t* d_offsets,
                                     const uint8_t* d_chars, size_t n, uint64_t* d_hash_out,
                                     uint32_t* d_rowid_out) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  // hash into a scratch column, sort from there into the output columns
  uint64_t* d_tmp = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_HASH_TMP0, n, &d_tmp));
  HMJ_TRY(hmj_hash_strings(ctx, d_offsets, d_chars, n, d_tmp));
  return hmj_impl_sort(ctx, d_tmp, nullptr, n, d_hash_out, d_rowid_out);
}

// ---------------------------------------------------------------- K8: top-bits partition
// Stands in for partition_only's counting loop and scatter (/root/reference
// partitioned_hash.h:57-60,75-79): partition p = hash >> (64 - bits).  The scatter reuses the
// onesweep pass with the digit taken from the top bits, so a partition is written as one
// contiguous, stable run (the reference's push_back order under a mutex is nondeterministic).
namespace {

constexpr int PART_THREADS = 512;

// counts[p] += 1 for p = key >> shift; shared-memory bins when they fit (bits <= 12)
template <typename CountT>
__global__ void __launch_bounds__(PART_THREADS)
partition_count_kernel(const uint64_t* __restrict__ keys, size_t n, int shift, int bins,
                       CountT* __restrict__ g_counts, int use_smem) {
  extern __shared__ uint32_t s_bins[];
  if (use_smem) {
    for (int i = threadIdx.x; i < bins; i += PART_THREADS) s_bins[i] = 0;
    __syncthreads();
  }
  const size_t stride = (size_t)gridDim.x * PART_THREADS;
  for (size_t i = (size_t)blockIdx.x * PART_THREADS + threadIdx.x; i < n; i += stride) {
    const uint32_t p = (uint32_t)(__ldcs(keys + i) >> shift);
    if (use_smem) atomicAdd(&s_bins[p], 1u);
    else atomicAdd(g_counts + p, (CountT)1);
  }
  if (use_smem) {
    __syncthreads();
    for (int i = threadIdx.x; i < bins; i += PART_THREADS) {
      const uint32_t c = s_bins[i];
      if (c) atomicAdd(g_counts + i, (CountT)c);
    }
  }
}

__global__ void partition_offsets_kernel(const uint32_t* __restrict__ excl, int bins, uint32_t n,
                                         uint64_t* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < bins) out[i] = excl[i];
  if (i == bins) out[i] = n;
}

}  // namespace

extern "C" int hmj_partition_counts(hmj_ctx* ctx, const uint64_t* d_hash, size_t n, int bits,
                                    uint64_t* d_counts) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, bits >= 1 && bits <= 16 && d_counts && (n == 0 || d_hash));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  const int bins = 1 << bits;
  HMJ_TRY(hmj_fill_async(ctx, d_counts, 0, (size_t)bins * 8, ctx->stream));
  if (n == 0) return HMJ_OK;
  const int use_smem = bits <= 12;
  const int grid = (int)min(hmj_div_up(n, PART_THREADS * 8), (size_t)ctx->num_sms * 4);
  partition_count_kernel<unsigned long long><<<grid, PART_THREADS, use_smem ? bins * 4 : 0, ctx->stream>>>(
      d_hash, n, 64 - bits, bins, reinterpret_cast<unsigned long long*>(d_counts), use_smem);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

extern "C" int hmj_partition_pairs(hmj_ctx* ctx, const uint64_t* d_hash, const uint32_t* d_rowid,
                                   size_t n, int bits, uint64_t* d_hash_out, uint32_t* d_rowid_out,
                                   uint64_t* d_part_offsets) {
  return hmj_partition_pairs_base(ctx, d_hash, d_rowid, 0, n, bits, d_hash_out, d_rowid_out,
                                  d_part_offsets);
}

extern "C" int hmj_partition_pairs_base(hmj_ctx* ctx, const uint64_t* d_hash,
                                        const uint32_t* d_rowid, uint32_t rowid_base, size_t n,
                                        int bits, uint64_t* d_hash_out, uint32_t* d_rowid_out,
                                        uint64_t* d_part_offsets) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, bits >= 1 && bits <= 8 && d_part_offsets);
  HMJ_ARG(ctx, n < (size_t(1) << 32) - (size_t(1) << 16));
  HMJ_ARG(ctx, n == 0 || (d_hash && d_hash_out && d_rowid_out));
  HMJ_ARG(ctx, d_hash != d_hash_out && (d_rowid == nullptr || d_rowid != d_rowid_out));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const int bins = 1 << bits;
  const uint32_t n32 = (uint32_t)n;
  uint32_t* d_hist = nullptr;
  const size_t hist_words = NUM_PASSES * RADIX + 2 * NUM_PASSES;
  HMJ_TRY(hmj_ws(ctx, WS_HIST, hist_words, &d_hist));
  uint32_t* d_skip = d_hist + NUM_PASSES * RADIX;
  uint32_t* d_counters = d_skip + NUM_PASSES;
  HMJ_TRY(hmj_fill_async(ctx, d_hist, 0, hist_words * sizeof(uint32_t), st));
  if (n) {
    const int grid = (int)min(hmj_div_up(n, PART_THREADS * 8), (size_t)ctx->num_sms * 4);
    partition_count_kernel<uint32_t><<<grid, PART_THREADS, RADIX * 4, st>>>(d_hash, n, 64 - bits, RADIX, d_hist, 1);
    HMJ_LAUNCHED(ctx);
  }
  radix_scan_kernel<<<1, RADIX, 0, st>>>(d_hist, n32, d_skip);
  HMJ_LAUNCHED(ctx);
  partition_offsets_kernel<<<(bins + 1 + 255) / 256, 256, 0, st>>>(d_hist, bins, n32, d_part_offsets);
  HMJ_LAUNCHED(ctx);
  if (n == 0) return HMJ_OK;
  const int tile_items = tile_items_for_cfg(0);
  const size_t tiles = hmj_div_up(n, tile_items);
  uint64_t* d_lb = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_LOOKBACK, tiles * RADIX, &d_lb));
  if (ctx->lb_seen_ptr != d_lb || ctx->lb_seen_bytes != ctx->ws[WS_LOOKBACK].bytes ||
      ctx->lb_epoch + NUM_PASSES >= 0xFFFF) {
    HMJ_TRY(hmj_fill_async(ctx, d_lb, 0, ctx->ws[WS_LOOKBACK].bytes, st));
    ctx->lb_seen_ptr = d_lb;
    ctx->lb_seen_bytes = ctx->ws[WS_LOOKBACK].bytes;
    ctx->lb_epoch = 0;
  }
  const uint32_t epoch = ++ctx->lb_epoch;
  return launch_pass<256, 16, 3>(ctx, d_hash, d_rowid, d_hash_out, d_rowid_out, n32, 64 - bits,
                                 d_hist, d_lb, d_counters, epoch, rowid_base);
}

// ---------------------------------------------------------------- partitioning without allocation
// hmj_partition_pairs_noalloc: the stable partition of hmj_partition_pairs_base with its scratch
// (digit histogram, tile counter, look-back descriptors) in a fixed array of device memory that
// the library carries itself, for callers that may not call cudaMalloc (code under CUDA graph
// capture, serving loops, harnesses that forbid it).  No counterpart in the reference.
// The existing entry points are untouched.
//   * The array is per GPU (every GPU that loads the library holds its NOALLOC_BYTES until the
//     process exits) and shared by every call on that GPU: it keeps the last call's contents, so
//     each call clears the words that must start at zero (histogram, counter, the descriptors of
//     the tiles it will use), in stream order before the kernels that use them.  Calls that
//     overlap on one GPU (two contexts / two streams) race on it: serialise them.
//   * The first call per GPU resolves the array's device address (cudaGetSymbolAddress -- the
//     name `g_noalloc_scratch` in host code is only a stand-in for it); make that call (n = 0 is
//     enough) before a region that must not touch the driver's allocator, e.g. a graph capture.
//   * No cudaMalloc / cudaFree / workspace growth, no profiling events: kernels only.
namespace {
constexpr size_t NOALLOC_MAX_ROWS = size_t(1) << 24;  // == HMJ_NOALLOC_MAX_ROWS (include/hmj.h)
constexpr size_t NOALLOC_TILE = OnesweepCfg<256, 16>::TILE;
constexpr size_t NOALLOC_MAX_TILES = (NOALLOC_MAX_ROWS + NOALLOC_TILE - 1) / NOALLOC_TILE;
constexpr size_t NOALLOC_HIST_WORDS = NUM_PASSES * RADIX + 2 * NUM_PASSES;
constexpr size_t NOALLOC_HIST_BYTES = (NOALLOC_HIST_WORDS * 4 + 255) & ~size_t(255);
constexpr size_t NOALLOC_LB_BYTES = NOALLOC_MAX_TILES * RADIX * 8;
constexpr size_t NOALLOC_BYTES = NOALLOC_HIST_BYTES + NOALLOC_LB_BYTES;  // 8 MiB + 8.25 KiB
static_assert(NOALLOC_MAX_ROWS == HMJ_NOALLOC_MAX_ROWS, "header and kernel file disagree on the limit");

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of the array, per GPU

// device address of this GPU's copy of the array (the device must be current)
int noalloc_scratch(hmj_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
    return hmj_fail(ctx, HMJ_EARG, "hmj_partition_pairs_noalloc: device ordinal beyond 63", __FILE__, __LINE__);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[ctx->device]) {
    void* p = nullptr;
    HMJ_CUDA(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255) != 0)
      return hmj_fail(ctx, HMJ_ECUDA, "hmj_partition_pairs_noalloc: no usable address for the scratch array", __FILE__, __LINE__);
    g_noalloc_ptr[ctx->device] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[ctx->device];
  return HMJ_OK;
}
}  // namespace

extern "C" int hmj_partition_pairs_noalloc(hmj_ctx* ctx, const uint64_t* d_hash,
                                           const uint32_t* d_rowid, uint32_t rowid_base, size_t n,
                                           int bits, uint64_t* d_hash_out, uint32_t* d_rowid_out,
                                           uint64_t* d_part_offsets) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, bits >= 1 && bits <= 8 && d_part_offsets);
  if (n > NOALLOC_MAX_ROWS)
    return hmj_fail(ctx, HMJ_EARG,
                    "hmj_partition_pairs_noalloc: n exceeds HMJ_NOALLOC_MAX_ROWS (16777216 rows): the fixed "
                    "scratch array holds no more; use hmj_partition_pairs_base",
                    __FILE__, __LINE__);
  HMJ_ARG(ctx, n == 0 || (d_hash && d_hash_out && d_rowid_out));
  HMJ_ARG(ctx, n == 0 || (d_hash != d_hash_out && (d_rowid == nullptr || d_rowid != d_rowid_out)));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  unsigned char* scratch = nullptr;
  HMJ_TRY(noalloc_scratch(ctx, &scratch));
  cudaStream_t st = ctx->stream;
  const int bins = 1 << bits;
  const uint32_t n32 = (uint32_t)n;
  uint32_t* d_hist = reinterpret_cast<uint32_t*>(scratch);
  uint32_t* d_skip = d_hist + NUM_PASSES * RADIX;
  uint32_t* d_counters = d_skip + NUM_PASSES;
  uint64_t* d_lb = reinterpret_cast<uint64_t*>(scratch + NOALLOC_HIST_BYTES);
  const size_t tiles = hmj_div_up(n, NOALLOC_TILE);
  // what the last call left behind and this one needs at zero
  HMJ_TRY(hmj_fill_async(ctx, d_hist, 0, NOALLOC_HIST_WORDS * sizeof(uint32_t), st));
  HMJ_TRY(hmj_fill_async(ctx, d_lb, 0, tiles * RADIX * 8, st));
  if (n) {
    const int grid = (int)min(hmj_div_up(n, PART_THREADS * 8), (size_t)ctx->num_sms * 4);
    partition_count_kernel<uint32_t><<<grid, PART_THREADS, RADIX * 4, st>>>(d_hash, n, 64 - bits, RADIX, d_hist, 1);
    HMJ_LAUNCHED(ctx);
  }
  radix_scan_kernel<<<1, RADIX, 0, st>>>(d_hist, n32, d_skip);
  HMJ_LAUNCHED(ctx);
  partition_offsets_kernel<<<(bins + 1 + 255) / 256, 256, 0, st>>>(d_hist, bins, n32, d_part_offsets);
  HMJ_LAUNCHED(ctx);
  if (n == 0) return HMJ_OK;
  using Cfg = OnesweepCfg<256, 16>;
  HMJ_CUDA(ctx, cudaFuncSetAttribute(onesweep_pass_kernel<256, 16, 3, 0, LB_WINDOW_DEFAULT>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM_BYTES));
  onesweep_pass_kernel<256, 16, 3, 0, LB_WINDOW_DEFAULT><<<(unsigned)tiles, 256, Cfg::SMEM_BYTES, st>>>(
      d_hash, d_rowid, d_hash_out, d_rowid_out, n32, 64 - bits, d_hist, d_lb, d_counters, 1u, rowid_base);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}
