// bucket_scatter.cu -- K3u: the "uniform hash" sort of (u64 hash, u32 row-id) pairs.
//
// The reference partitions on the top hash bits first (radix_hash.h:351-423) and then refines
// every partition independently (radix_hash.h:122-292): an MSD radix sort, which needs no stable
// passes because a partition is never mixed with another again.  On the GPU that removes the
// decoupled look-back chain of the stable LSD pass (profiles/r1c: a tile of the onesweep pass
// spends 6 of its 14 us walking back through ~40 earlier tiles) and the warp ranking:
//   level 1/2  bucket_scatter_kernel: a CTA counts its tile's digits with shared-memory atomics,
//              reserves room in every destination bucket with ONE global atomicAdd per bucket
//              (order of arrival, not of tile id) and writes its runs there.  Buckets live in
//              fixed-capacity regions (mean + 8 sigma of a uniform hash), so no histogram pass
//              is needed before the data moves.
//   finish     segment_sort_kernel (segment_sort.cu) sorts every final bucket (~2^11..2^12
//              pairs) in shared memory on the remaining key bits, breaks ties on the row id, and
//              writes the dense, fully ordered output.  (key, row id) ascending is exactly what
//              the stable sort produces from iota row ids, so the result is canonical.
// A bucket that overflows its region (skewed or heavily duplicated hashes) raises *flag; the
// caller then runs the histogram-planned stable sort on the untouched input.
// Algorithmic bytes per pair: level 1 reads 8 (+4 with row ids) and writes 12; level 2 reads 12
// and writes 12; finish reads 12 and writes 12.
#include <math.h>

#include <algorithm>

#include "common.cuh"

namespace {

constexpr int US_THREADS = 256;
constexpr int US_ITEMS = 16;
constexpr int US_TILE = US_THREADS * US_ITEMS;
constexpr int US_MAX_BINS = 512;  // 9-bit levels run with the 1024-thread configuration only
constexpr int US_CURSOR_STRIDE = 1;  // words between cursors (see uniform_plan)
constexpr size_t US_SMEM = (size_t)US_TILE * 12 + 3 * US_MAX_BINS * 4 + 64;

// Input: `regions_in` regions of `in_stride` slots, region r holding in_counts[r] pairs (or one
// region of n_single pairs when in_counts == nullptr).  CTA b takes tile b % tiles_per_region of
// region b / tiles_per_region.  Output bucket of a pair = (r << bits) | digit, stored in the
// region [bucket * out_cap, (bucket + 1) * out_cap) at the slot reserved through cursors[bucket].
__global__ void __launch_bounds__(US_THREADS, 3)
bucket_scatter_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                      const uint32_t* __restrict__ in_counts, uint32_t in_stride,
                      uint32_t n_single, uint32_t tiles_per_region, uint32_t iota_base,
                      uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int shift,
                      int bits, uint32_t out_cap, uint32_t* __restrict__ cursors,
                      uint32_t* __restrict__ flag) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem_raw);
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys + US_TILE);
  uint32_t* s_cnt = s_vals + US_TILE;       // [bins] tile counts
  uint32_t* s_off = s_cnt + US_MAX_BINS;    // [bins] start of the digit inside the tile
  uint32_t* s_gbase = s_off + US_MAX_BINS;  // [bins] global slot of the digit's run minus s_off
  uint32_t* s_scan = s_gbase + US_MAX_BINS; // [8]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t r = blockIdx.x / tiles_per_region, j = blockIdx.x % tiles_per_region;
  // (a region that overflowed holds in_stride pairs at most; the flag is up and the result void)
  const uint32_t count = in_counts ? min(in_counts[r * US_CURSOR_STRIDE], in_stride) : n_single;
  const uint64_t tile_first = (uint64_t)j * US_TILE;
  if (tile_first >= count) return;
  // the attempt is already void once any bucket has overflowed
  if (__syncthreads_or(tid == 0 && *reinterpret_cast<volatile uint32_t*>(flag) != 0)) return;
  const uint32_t valid = (uint32_t)min((uint64_t)US_TILE, count - tile_first);
  const uint32_t in_base = r * in_stride + (uint32_t)tile_first;
  const int bins = 1 << bits;
  const uint32_t mask = (uint32_t)bins - 1;

  if (tid < bins) s_cnt[tid] = 0;
  uint64_t key[US_ITEMS];
#pragma unroll
  for (int k = 0; k < US_ITEMS; ++k) {
    const uint32_t i = tid + k * US_THREADS;
    key[k] = i < valid ? __ldcs(keys_in + in_base + i) : 0;
  }
  __syncthreads();

  uint16_t slot[US_ITEMS];
#pragma unroll
  for (int k = 0; k < US_ITEMS; ++k) {
    const uint32_t i = tid + k * US_THREADS;
    if (i < valid) slot[k] = (uint16_t)atomicAdd(&s_cnt[(uint32_t)(key[k] >> shift) & mask], 1u);
  }
  __syncthreads();

  // one reservation per destination bucket; the scan over the tile counts runs while the global
  // atomics are in flight
  uint32_t c = 0, old = 0, bucket = 0;
  if (tid < bins) {
    c = s_cnt[tid];
    bucket = (r << bits) | (uint32_t)tid;
    if (c) old = atomicAdd(&cursors[(size_t)bucket * US_CURSOR_STRIDE], c);
  }
  {
    uint32_t x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_scan[warp] = x;
    __syncthreads();
    uint32_t wt = lane < warp ? s_scan[lane] : 0;  // 8 warps
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) wt += __shfl_xor_sync(0xffffffffu, wt, o);
    wt = __shfl_sync(0xffffffffu, wt, 0);
    if (tid < bins) s_off[tid] = wt + x - c;
  }
  __syncthreads();

  // tile-sorted order in shared memory, so that a bucket's run leaves as consecutive stores
  uint16_t pos[US_ITEMS];
#pragma unroll
  for (int k = 0; k < US_ITEMS; ++k) {
    const uint32_t i = tid + k * US_THREADS;
    if (i < valid) {
      const uint32_t p = s_off[(uint32_t)(key[k] >> shift) & mask] + slot[k];
      pos[k] = (uint16_t)p;
      s_keys[p] = key[k];
    }
  }
#pragma unroll
  for (int k = 0; k < US_ITEMS; ++k) {
    const uint32_t i = tid + k * US_THREADS;
    if (i < valid) s_vals[pos[k]] = vals_in ? __ldcs(vals_in + in_base + i) : in_base + i + iota_base;
  }
  bool over = false;
  if (tid < bins) {
    over = c != 0 && (old > out_cap || c > out_cap - old);
    s_gbase[tid] = bucket * out_cap + old - s_off[tid];
  }
  if (__syncthreads_or(over)) {
    if (tid == 0) atomicOr(flag, 1u);
    return;
  }
#pragma unroll 4
  for (uint32_t i = tid; i < valid; i += US_THREADS) {
    const uint64_t kk = s_keys[i];
    const uint32_t dst = s_gbase[(uint32_t)(kk >> shift) & mask] + i;
    keys_out[dst] = kk;
    vals_out[dst] = s_vals[i];
  }
}

// ---- TMA-pipelined variant -------------------------------------------------------------
// Same algorithm, persistent CTAs (two per SM).  A tile's key and row-id columns arrive in shared
// memory by 1-D bulk async copies (cp.async.bulk, completion on an mbarrier); the copy of the
// NEXT tile is issued before the current one is processed, so ~96 KB of loads per SM are always
// in flight and no thread holds registers for them (the plain kernel above exposes the row-id
// load latency: profiles/r1d, 54 % of its stall samples).  The tile is permuted in place in its
// buffer and leaves with coalesced stores.
constexpr int UT_ITEMS = 8;
// (Measured and removed in round 2: ranking a tile's items with a warp-level multisplit --
// match.any, or one ballot per digit bit -- on warp-private u16 counters instead of one
// shared-memory atomicAdd per item: 2.66 ms / 2.33 ms per 256M-pair launch against 1.35 ms.  The
// dependent load-store chain per item costs more than the atomics' LSU time.)
template <int THREADS>
struct UtCfg {
  static constexpr int TILE = THREADS * UT_ITEMS;
  static constexpr int BINS = THREADS > 512 ? 512 : 256;   // most bins a level of this shape cuts
  static constexpr int SLACK = 64;  // slots behind the tile: room to start runs at padded positions (DIST)
  static constexpr size_t BUF_BYTES = (size_t)(TILE + SLACK) * 12;
  // buffers, [cnt][off][gbase] per bin, region ends and starts (<= 512 regions), scan, bars, peers
  static constexpr size_t SMEM = 2 * BUF_BYTES + 3 * BINS * 4 + 2 * 512 * 4 + 32 * 4 + 64 + 16 * 8;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src, uint32_t bytes,
                                          uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// DIST: the destination regions are spread over the GPUs of the job (multi-GPU level 1, the
// exchange step of SURVEY.md 8e fused into the scatter): region g of the 2^bits belongs to GPU
// g >> own_bits and lives in that GPU's memory, reached through peer pointers (NVLink) with plain
// stores.  Every source GPU has its own sub-region in every region, so the slot reservation is
// an atomicAdd on a LOCAL cursor (remote same-address atomics from all GPUs on one cursor
// serialise at ~30 ns each and bound the first version of this kernel at 240 GB/s of NVLink).
// region_shift (level 2 of such a job): input region t / tiles_per_region is a sub-region; the
// bucket prefix is the region it belongs to, (t / tiles_per_region) >> region_shift.
// in_starts (the multi-GPU join takes a relation in parts): input region r holds the pairs
// [in_starts[r], in_counts[r]); earlier parts of the region were processed by an earlier launch.
// The first tile of a region then starts at in_starts[r] rounded down to 4 pairs (bulk copies need
// 16-byte aligned sources) and ignores the `lo` pairs before the start.
// VARIANT compiles the rarely used paths out of the common kernel: 0 plain, 1 region starts (local
// levels of the multi-GPU join taken in parts), 2 prefix check (first level of a sample-derived plan)
template <int UT_THREADS, int MIN_BLOCKS, bool DIST, int VARIANT>
__global__ void __launch_bounds__(UT_THREADS, MIN_BLOCKS)
bucket_scatter_tma_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                          const uint32_t* __restrict__ in_counts, const uint32_t* __restrict__ in_starts,
                          uint32_t in_stride, uint32_t n_single, uint32_t tiles_per_region,
                          uint32_t total_tiles, uint32_t iota_base, uint64_t* __restrict__ keys_out,
                          uint32_t* __restrict__ vals_out, int shift, int bits, uint32_t out_cap,
                          uint32_t* __restrict__ cursors, uint32_t* __restrict__ flag,
                          uint32_t* __restrict__ tile_counter, const hmj_peer_regions peers,
                          int region_shift, uint64_t prefix, int prefix_bits) {
  using Cfg = UtCfg<UT_THREADS>;
  constexpr bool STARTS = VARIANT == 1, PREFIX = VARIANT == 2;
  constexpr int UT_TILE = Cfg::TILE;
  constexpr size_t UT_BUF_BYTES = Cfg::BUF_BYTES;
  extern __shared__ __align__(128) unsigned char smem_tma[];
  uint32_t* s_cnt = reinterpret_cast<uint32_t*>(smem_tma + 2 * UT_BUF_BYTES);
  uint32_t* s_off = s_cnt + Cfg::BINS;
  uint32_t* s_gbase = s_off + Cfg::BINS;
  uint32_t* s_rcount = s_gbase + Cfg::BINS;    // [regions <= 512] end of the pairs of an input region
  uint32_t* s_rstart = s_rcount + 512;         // [regions <= 512] ... and their start
  uint32_t* s_scan = s_rstart + 512;           // [32]
  uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_scan + 32);  // [2]
  uint64_t** s_pk = reinterpret_cast<uint64_t**>(s_bar + 2);   // [8] peer key columns (DIST)
  uint32_t** s_pv = reinterpret_cast<uint32_t**>(s_pk + 8);    // [8] peer row-id columns
  uint32_t* s_next = reinterpret_cast<uint32_t*>(s_pv + 8);    // [2] tile ids handed to the CTA

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // Tiles are handed out by a global counter, not by block index: a CTA that starts late (the
  // kernel shares the SMs with another stream's kernel in the multi-GPU join) simply takes fewer
  // tiles.  The id of the tile after next is fetched one iteration ahead, so the atomic's round
  // trip never shows.
  if (tid == 0) {
    const uint32_t v = atomicAdd(tile_counter, 2u);
    s_next[0] = v;
    s_next[1] = v + 1;
  }
  if (DIST && tid < 8) {
    s_pk[tid] = peers.keys[tid];
    s_pv[tid] = peers.vals[tid];
  }
  const int bins = 1 << bits;
  const uint32_t mask = (uint32_t)bins - 1;
  if (tid == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (in_counts) {
    const uint32_t regions = (total_tiles + tiles_per_region - 1) / tiles_per_region;
    for (uint32_t r = tid; r < regions; r += UT_THREADS) {
      s_rcount[r] = min(in_counts[r * US_CURSOR_STRIDE], in_stride);
      if (STARTS) s_rstart[r] = min(in_starts[r * US_CURSOR_STRIDE], in_stride);
    }
  }
  __syncthreads();

  // tile t -> (first slot, pairs to ignore at its head, end of its pairs); 0 = nothing there
  auto tile_of = [&](uint32_t t, uint32_t* in_base, uint32_t* lo) -> uint32_t {
    const uint32_t r = t / tiles_per_region, j = t % tiles_per_region;
    const uint32_t count = in_counts ? s_rcount[r] : n_single;
    const uint32_t start = (STARTS && in_counts) ? s_rstart[r] : 0u;
    const uint32_t aligned = start & ~3u;
    const uint64_t first = (uint64_t)aligned + (uint64_t)j * UT_TILE;
    if (first >= count) return 0;
    *in_base = r * in_stride + (uint32_t)first;
    *lo = (STARTS && j == 0) ? start - aligned : 0u;
    return (uint32_t)min((uint64_t)UT_TILE, count - first);
  };
  // Starts the copies of tile t into buffer b (full tiles only; a partial tile is loaded with
  // plain loads when its turn comes).  Every thread learns whether a copy is in flight, thread 0
  // issues it.
  auto issue = [&](uint32_t t, int b) -> bool {
    uint32_t in_base = 0, lo = 0;
    const uint32_t valid = t < total_tiles ? tile_of(t, &in_base, &lo) : 0;
    if (valid != (uint32_t)UT_TILE) return false;
    if (tid == 0) {
      unsigned char* buf = smem_tma + (size_t)b * UT_BUF_BYTES;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      const uint32_t bytes = UT_TILE * 8 + (vals_in ? UT_TILE * 4 : 0);
      mbar_expect_tx(&s_bar[b], bytes);
      bulk_load(buf, keys_in + in_base, UT_TILE * 8, &s_bar[b]);
      if (vals_in) bulk_load(buf + (UT_TILE + Cfg::SLACK) * 8, vals_in + in_base, UT_TILE * 4, &s_bar[b]);
    }
    return true;
  };

  uint32_t phase[2] = {0, 0};
  uint32_t t = s_next[0], t_next = s_next[1];
  __syncthreads();  // s_next is rewritten from the first iteration on
  bool pending_next = issue(t, 0);
  for (int it = 0; t < total_tiles; ++it) {
    const int b = it & 1;
    // another CTA's overflow ends the attempt.  The flag is read here and looked at only behind
    // the tile's arrival below, so its round trip to L2 hides behind the copy instead of holding
    // the whole CTA at a barrier of its own (16 % of the stall samples in profiles/r2s_scatter_digest.txt).
    uint32_t failed = 0;
    if (tid == 0) failed = *reinterpret_cast<volatile uint32_t*>(flag);
    const bool pending_cur = pending_next;
    pending_next = issue(t_next, b ^ 1);
    uint32_t fetched = 0;
    if (tid == 0) fetched = atomicAdd(tile_counter, 1u);
    uint32_t in_base = 0, lo_rt = 0;
    const uint32_t valid = tile_of(t, &in_base, &lo_rt);
    const uint32_t lo = STARTS ? lo_rt : 0u;
    if (valid <= lo) {
      if (tid == 0) s_next[b] = fetched;
      __syncthreads();
      t = t_next;
      t_next = s_next[b];
      continue;
    }
    uint64_t* b_keys = reinterpret_cast<uint64_t*>(smem_tma + (size_t)b * UT_BUF_BYTES);
    uint32_t* b_vals = reinterpret_cast<uint32_t*>(b_keys + UT_TILE + Cfg::SLACK);
    if (tid < bins) s_cnt[tid] = 0;
    if (pending_cur) {
      mbar_wait(&s_bar[b], phase[b]);
      phase[b] ^= 1;
    } else {
      for (uint32_t i = tid; i < valid; i += UT_THREADS) {
        b_keys[i] = keys_in[in_base + i];
        if (vals_in) b_vals[i] = vals_in[in_base + i];
      }
    }
    // (the counters of the previous tile were last read before its closing barrier)
    if (__syncthreads_or(failed != 0)) {
      // never leave with a bulk copy still writing into this CTA's shared memory
      if (pending_next) mbar_wait(&s_bar[b ^ 1], phase[b ^ 1]);
      return;
    }

    // prefix_bits > 0 (first level of a plan that skips the hash bits all rows were found to share
    // in a sample): a row with other top bits would land in a wrong bucket -- it voids the attempt
    bool stray = false;
    uint64_t key[UT_ITEMS];
    uint16_t slot[UT_ITEMS];
#pragma unroll
    for (int k = 0; k < UT_ITEMS; ++k) {
      const uint32_t i = tid + k * UT_THREADS;
      if ((!STARTS || i >= lo) && i < valid) {
        key[k] = b_keys[i];
        slot[k] = (uint16_t)atomicAdd(&s_cnt[(uint32_t)(key[k] >> shift) & mask], 1u);
        if (PREFIX) stray |= (key[k] >> (64 - prefix_bits)) != prefix;
      }
    }
    __syncthreads();

    uint32_t c = 0, old = 0, bucket = 0;
    if (tid < bins) {
      c = s_cnt[tid];
      bucket = (((t / tiles_per_region) >> region_shift) << bits) | (uint32_t)tid;
      if (c) old = atomicAdd(&cursors[(size_t)bucket * US_CURSOR_STRIDE], c);
      // DIST (level 1 of the job, bucket == tid): my sub-region of the owner's region
      if (DIST) bucket = (bucket & ((1u << peers.own_bits) - 1)) * (uint32_t)peers.world + (uint32_t)peers.rank;
    }
    uint32_t val[UT_ITEMS];
#pragma unroll
    for (int k = 0; k < UT_ITEMS; ++k) {
      const uint32_t i = tid + k * UT_THREADS;
      if ((!STARTS || i >= lo) && i < valid) val[k] = vals_in ? b_vals[i] : in_base + i + iota_base;
    }
    {
      uint32_t x = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      if (lane == 31) s_scan[warp] = x;
      __syncthreads();
      uint32_t wt = lane < warp ? s_scan[lane] : 0;  // <= 32 warps
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) wt += __shfl_xor_sync(0xffffffffu, wt, o);
      if (tid < bins) s_off[tid] = wt + x - c;
      if (DIST && tid < bins) s_gbase[tid] = old;  // (the bulk-store path below reads it back)
    }
    __syncthreads();  // s_off ready; every thread holds its keys and row ids in registers
    // DIST with few regions: a region's run leaves as ONE bulk async store per column
    // (cp.async.bulk shared -> peer global), which needs source and destination equally aligned
    // to 16 bytes.  The run of region d therefore starts at a tile position congruent to its
    // first global slot modulo 4 pairs (up to 3 slots of padding per region, in the slack behind
    // the tile).
    const bool bulk_out = DIST && peers.bulk && bins <= 16;
    if (bulk_out) {
      if (tid == 0) {
        uint32_t run = 0;
        for (int d = 0; d < bins; ++d) {
          run += (s_gbase[d] - run) & 3u;
          s_off[d] = run;
          run += s_cnt[d];
        }
      }
      __syncthreads();
    }

#pragma unroll
    for (int k = 0; k < UT_ITEMS; ++k) {
      const uint32_t i = tid + k * UT_THREADS;
      if ((!STARTS || i >= lo) && i < valid) {
        const uint32_t p = s_off[(uint32_t)(key[k] >> shift) & mask] + slot[k];
        b_keys[p] = key[k];
        b_vals[p] = val[k];
      }
    }
    bool over = stray;
    if (tid < bins) {
      over |= c != 0 && (old > out_cap || c > out_cap - old);
      if (!bulk_out) s_gbase[tid] = bucket * out_cap + old - s_off[tid];  // (bulk: keeps `old`)
    }
    if (__syncthreads_or(over)) {
      if (tid == 0) atomicOr(flag, 1u);
      if (pending_next) mbar_wait(&s_bar[b ^ 1], phase[b ^ 1]);
      return;
    }
    const uint32_t live = valid - lo;  // the permuted pairs sit at [0, live)
    if (bulk_out) {
      // warp d sends region d: head and tail pairs (to the 4-pair boundaries of the destination)
      // with plain stores, everything between with two bulk stores
      if (warp < bins) {
        const uint32_t d = (uint32_t)warp, cnt_d = s_cnt[d];
        if (cnt_d) {
          const uint32_t p0 = s_off[d], old_d = s_gbase[d];
          const uint32_t sub = (d & ((1u << peers.own_bits) - 1)) * (uint32_t)peers.world + (uint32_t)peers.rank;
          const uint32_t g0 = sub * out_cap + old_d;
          const uint32_t o = d >> peers.own_bits;
          const uint32_t head = min(cnt_d, (4u - (g0 & 3u)) & 3u);
          const uint32_t mid = (cnt_d - head) & ~3u;
          const uint32_t tail = cnt_d - head - mid;
          if ((uint32_t)lane < head) {
            s_pk[o][g0 + lane] = b_keys[p0 + lane];
            s_pv[o][g0 + lane] = b_vals[p0 + lane];
          }
          if ((uint32_t)lane < tail) {
            s_pk[o][g0 + head + mid + lane] = b_keys[p0 + head + mid + lane];
            s_pv[o][g0 + head + mid + lane] = b_vals[p0 + head + mid + lane];
          }
          if (lane == 0 && mid) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(s_pk[o] + g0 + head),
                         "r"(smem_u32(b_keys + p0 + head)), "r"(mid * 8u)
                         : "memory");
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(s_pv[o] + g0 + head),
                         "r"(smem_u32(b_vals + p0 + head)), "r"(mid * 4u)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            // the buffer is refilled after the barrier below: its bytes must have been read by then
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          }
        }
      }
    } else
#pragma unroll 4
    for (uint32_t i = tid; i < live; i += UT_THREADS) {
      const uint64_t kk = b_keys[i];
      const uint32_t d = (uint32_t)(kk >> shift) & mask;
      const uint32_t dst = s_gbase[d] + i;
      if (DIST) {
        const uint32_t o = d >> peers.own_bits;
        s_pk[o][dst] = kk;
        s_pv[o][dst] = b_vals[i];
      } else {
        keys_out[dst] = kk;
        vals_out[dst] = b_vals[i];
      }
    }
    if (tid == 0) s_next[b] = fetched;
    __syncthreads();  // the buffer may be refilled from here on
    t = t_next;
    t_next = s_next[b];
  }
}

// after a DIST level 1: this source's cursor of region r goes to the owner's count table `part`
__global__ void push_counts_kernel(const uint32_t* __restrict__ cursors, uint32_t regions,
                                   const hmj_peer_regions peers, uint32_t table_offset) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= regions) return;
  const uint32_t local = r & ((1u << peers.own_bits) - 1);
  peers.counts[r >> peers.own_bits][table_offset + local * (uint32_t)peers.world + (uint32_t)peers.rank] = cursors[r];
}

// exclusive scan of the final bucket counts -> dense output offsets (one CTA; <= 2^18 buckets)
__global__ void __launch_bounds__(1024)
bucket_offsets_kernel(const uint32_t* __restrict__ counts, uint32_t buckets,
                      uint32_t* __restrict__ offsets) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < buckets; base += 1024) {
    const uint32_t i = base + tid;
    const uint32_t c = i < buckets ? counts[(size_t)i * US_CURSOR_STRIDE] : 0;
    uint32_t x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    uint32_t wt = lane < warp ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wt += __shfl_xor_sync(0xffffffffu, wt, o);
    const uint32_t carry = s_carry;
    if (i < buckets) offsets[i] = carry + wt + x - c;
    __syncthreads();
    if (tid == 1023) s_carry = carry + wt + x;
    __syncthreads();
  }
  if (tid == 0) offsets[buckets] = s_carry;
}

}  // namespace

namespace {

int uniform_from_env() {
  const char* e = getenv("HMJ_SORT_UNIFORM");
  return e ? atoi(e) : 1;
}

struct ScatterLaunch {
  const uint64_t* keys_in; const uint32_t* vals_in; const uint32_t* in_counts;
  const uint32_t* in_starts;                // != nullptr: region r holds [in_starts[r], in_counts[r])
  uint32_t in_stride, n_single, tiles_per_region, total_tiles, iota_base;
  uint64_t* keys_out; uint32_t* vals_out;
  int shift, bits;
  uint32_t out_cap;
  uint32_t *cursors, *flag, *tile_counter;  // tile_counter: zeroed by the caller
  const hmj_peer_regions* peers;            // != nullptr: regions on the GPUs of the job (level 1)
  int region_shift;
  uint64_t prefix;                          // prefix_bits > 0: every key's top bits must equal prefix
  int prefix_bits;
};

template <int T, int B, bool DIST, int VARIANT>
int launch_tma(hmj_ctx* ctx, const ScatterLaunch& a, uint32_t ctas_per_sm, cudaStream_t st) {
  using Cfg = UtCfg<T>;
  static_assert(Cfg::SMEM * B + 1024 * B <= 233472, "scatter configuration does not fit an SM");
  auto* kern = bucket_scatter_tma_kernel<T, B, DIST, VARIANT>;
  HMJ_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM));
  const uint32_t per_sm = ctas_per_sm == 0 || ctas_per_sm > (uint32_t)B ? (uint32_t)B : ctas_per_sm;
  const uint32_t grid = std::min(a.total_tiles, (uint32_t)ctx->num_sms * per_sm);
  kern<<<grid, T, Cfg::SMEM, st>>>(a.keys_in, a.vals_in, a.in_counts, a.in_starts, a.in_stride,
                                   a.n_single, a.tiles_per_region, a.total_tiles, a.iota_base,
                                   a.keys_out, a.vals_out, a.shift, a.bits, a.out_cap, a.cursors,
                                   a.flag, a.tile_counter, a.peers ? *a.peers : hmj_peer_regions(),
                                   a.region_shift, a.prefix, a.prefix_bits);
  HMJ_LAUNCHED(ctx);
  ctx->tm.scatter_levels++;
  return HMJ_OK;
}

// pairs per tile of the configuration a level of `bits` bits runs with
uint32_t tile_items_for(int bits) { return bits > 8 ? 8192u : 4096u; }

// One bucket-scatter launch (TMA-pipelined kernel).  ctas_per_sm = 0: as many as fit.
int launch_scatter(hmj_ctx* ctx, const ScatterLaunch& a, uint32_t ctas_per_sm, cudaStream_t st) {
  if (a.total_tiles == 0) return HMJ_OK;
  const bool dist = a.peers != nullptr;
  const int variant = a.in_starts != nullptr ? 1 : a.prefix_bits > 0 ? 2 : 0;
  if (dist) {
    HMJ_ARG(ctx, variant == 0);
    return a.bits > 8 ? launch_tma<1024, 1, true, 0>(ctx, a, ctas_per_sm, st)
                      : launch_tma<512, 2, true, 0>(ctx, a, ctas_per_sm, st);
  }
  if (a.bits > 8) {  // 512 bins: 1024-thread tiles of 8192 pairs
    return variant == 1 ? launch_tma<1024, 1, false, 1>(ctx, a, ctas_per_sm, st)
         : variant == 2 ? launch_tma<1024, 1, false, 2>(ctx, a, ctas_per_sm, st)
                        : launch_tma<1024, 1, false, 0>(ctx, a, ctas_per_sm, st);
  }
  return variant == 1 ? launch_tma<512, 2, false, 1>(ctx, a, ctas_per_sm, st)
       : variant == 2 ? launch_tma<512, 2, false, 2>(ctx, a, ctas_per_sm, st)
                      : launch_tma<512, 2, false, 0>(ctx, a, ctas_per_sm, st);
}

}  // namespace

// Final buckets of at most 4096 pairs on average for `n_plan` rows (the larger relation of a
// join, so that both sides share the bucket numbering); one or two scatter levels of <= 8 bits.
hmj_uniform_plan hmj_uniform_plan_for(size_t n_plan, int skip_bits) {
  hmj_uniform_plan p = {};
  const uint32_t cap_seg = (uint32_t)hmj_segment_sort_cap();
  if (uniform_from_env() == 0 || n_plan <= cap_seg) return p;
  int qb = 0;
  while (((n_plan + ((size_t)1 << qb) - 1) >> qb) > 4096) ++qb;
  if (qb > 18 || skip_bits < 0 || skip_bits + qb + 12 > 64) return p;
  p.qb = qb;
  p.skip = skip_bits;
  p.cap_seg = cap_seg;
  if (qb <= 8) {
    p.levels = 1;
    p.bits[0] = qb;
  } else {
    p.levels = 2;
    p.bits[0] = (qb + 1) / 2;
    p.bits[1] = qb - p.bits[0];
    // HMJ_UNIFORM_BITS0: width of the first level (scripts/sweep_bits.py, the counterpart of the
    // reference's find_k_bench.cc:116-129 sweep of the partition width)
    if (const char* e = getenv("HMJ_UNIFORM_BITS0")) {
      const int v = atoi(e);
      if (v >= 1 && v <= 9 && qb - v >= 1 && qb - v <= 9) {
        p.bits[0] = v;
        p.bits[1] = qb - v;
      }
    }
  }
  return p;
}

// Scatters (keys_in[i], vals_in ? vals_in[i] : iota_base + i) into the 2^qb padded buckets of
// plan `p` (bucket = the qb hash bits below the p.skip topmost ones).  Bucket b ends
// up at out->keys/vals[b * cap_seg ...] with out->counts[b] pairs; *out->flag != 0 (device)
// means a region overflowed and the buckets are void.  `set` selects the workspace slots so
// that two relations can be held at once (0 or 1).  Nothing is synchronised.
int hmj_uniform_scatter(hmj_ctx* ctx, const uint64_t* keys_in, const uint32_t* vals_in, size_t n,
                        uint32_t iota_base, const hmj_uniform_plan& p, int set,
                        hmj_uniform_buckets* out) {
  cudaStream_t st = ctx->stream;
  const uint32_t n32 = (uint32_t)n;
  const uint32_t regions1 = 1u << p.bits[0], buckets = 1u << p.qb;
  uint32_t cap1 = 0;
  if (p.levels == 2) {
    const double avg1 = (double)n * (p.density > 1.0 ? p.density : 1.0) / (double)regions1;
    cap1 = ((uint32_t)(avg1 + 8.0 * sqrt(avg1) + 64.0) + 255u) & ~255u;
  }
  // scratch words: [flag, tile counters: 8 words][level-1 cursors][final cursors][final offsets + 1]
  uint32_t* d_words = nullptr;
  const size_t words = 8 + (size_t)regions1 + buckets + buckets + 1;
  HMJ_TRY(hmj_ws(ctx, set ? WS_US_CURSORS_1 : WS_US_CURSORS, words, &d_words));
  uint32_t* d_flag = d_words;
  uint32_t* d_cur1 = d_words + 8;
  uint32_t* d_cur2 = p.levels == 2 ? d_cur1 + regions1 : d_cur1;
  HMJ_TRY(hmj_fill_async(ctx, d_words, 0, (8 + (size_t)regions1 + buckets) * sizeof(uint32_t), st));

  uint64_t *d_k1 = nullptr, *d_k2 = nullptr;
  uint32_t *d_v1 = nullptr, *d_v2 = nullptr;
  const size_t slots2 = (size_t)buckets * p.cap_seg;
  HMJ_TRY(hmj_ws(ctx, set ? WS_SORT_KEYS_C : WS_SORT_KEYS_B, slots2, &d_k2));
  HMJ_TRY(hmj_ws(ctx, set ? WS_SORT_VALS_C : WS_SORT_VALS_B, slots2, &d_v2));
  if (p.levels == 2) {
    const size_t slots1 = (size_t)regions1 * cap1;
    HMJ_TRY(hmj_ws(ctx, WS_SORT_KEYS_A, slots1, &d_k1));
    HMJ_TRY(hmj_ws(ctx, WS_SORT_VALS_A, slots1, &d_v1));
  }
  // bulk copies need 16-byte aligned sources (tile starts are multiples of 4096 pairs)
  const bool tma = uniform_from_env() != 2 && (reinterpret_cast<uintptr_t>(keys_in) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(vals_in) & 15) == 0;
  const bool wide = p.bits[0] > 8 || p.bits[1] > 8;  // 512 bins: 1024-thread tiles of 8192 pairs
  HMJ_ARG(ctx, tma || !wide);
  cudaEvent_t ev;
  // level 1: the raw key column -> 2^bits[0] regions
  uint64_t* k_out1 = p.levels == 2 ? d_k1 : d_k2;
  uint32_t* v_out1 = p.levels == 2 ? d_v1 : d_v2;
  const uint32_t cap_out1 = p.levels == 2 ? cap1 : p.cap_seg;
  if (n) {
    HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_SCATTER, (double)n * (vals_in ? 24.0 : 20.0), &ev));
    if (tma) {
      ScatterLaunch a = {};
      a.keys_in = keys_in; a.vals_in = vals_in; a.n_single = n32; a.iota_base = iota_base;
      a.total_tiles = (uint32_t)hmj_div_up(n, tile_items_for(p.bits[0]));
      a.tiles_per_region = a.total_tiles;
      a.keys_out = k_out1; a.vals_out = v_out1;
      a.shift = 64 - p.skip - p.bits[0]; a.bits = p.bits[0]; a.out_cap = cap_out1;
      a.cursors = d_cur1; a.flag = d_flag; a.tile_counter = d_words + 1;
      a.prefix = p.prefix; a.prefix_bits = p.check_prefix ? p.skip : 0;
      HMJ_TRY(launch_scatter(ctx, a, 0, st));
    } else {
      HMJ_ARG(ctx, !p.check_prefix);
      HMJ_CUDA(ctx, cudaFuncSetAttribute(bucket_scatter_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)US_SMEM));
      const uint32_t tiles = (uint32_t)hmj_div_up(n, US_TILE);
      bucket_scatter_kernel<<<tiles, US_THREADS, US_SMEM, st>>>(
          keys_in, vals_in, nullptr, 0, n32, tiles, iota_base, k_out1, v_out1,
          64 - p.skip - p.bits[0], p.bits[0], cap_out1, d_cur1, d_flag);
      HMJ_LAUNCHED(ctx);
      ctx->tm.scatter_levels++;
    }
    hmj_prof_end(ctx, ev);
  }
  if (p.levels == 2 && n) {
    HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_SCATTER, (double)n * 24.0, &ev));
    if (tma) {
      ScatterLaunch a = {};
      a.keys_in = d_k1; a.vals_in = d_v1; a.in_counts = d_cur1; a.in_stride = cap1;
      a.tiles_per_region = (uint32_t)hmj_div_up(cap1, tile_items_for(p.bits[1]));
      a.total_tiles = regions1 * a.tiles_per_region;
      a.keys_out = d_k2; a.vals_out = d_v2;
      a.shift = 64 - p.skip - p.qb; a.bits = p.bits[1]; a.out_cap = p.cap_seg;
      a.cursors = d_cur2; a.flag = d_flag; a.tile_counter = d_words + 2;
      HMJ_TRY(launch_scatter(ctx, a, 0, st));
    } else {
      const uint32_t tpr = (uint32_t)hmj_div_up(cap1, US_TILE);
      bucket_scatter_kernel<<<regions1 * tpr, US_THREADS, US_SMEM, st>>>(
          d_k1, d_v1, d_cur1, cap1, 0, tpr, 0, d_k2, d_v2, 64 - p.skip - p.qb, p.bits[1],
          p.cap_seg, d_cur2, d_flag);
      HMJ_LAUNCHED(ctx);
      ctx->tm.scatter_levels++;
    }
    hmj_prof_end(ctx, ev);
  }
  out->keys = d_k2;
  out->vals = d_v2;
  out->counts = d_cur2;
  out->offsets = d_cur1 + regions1 + buckets;
  out->flag = d_flag;
  out->buckets = buckets;
  return HMJ_OK;
}

// Sorts (keys_in[i], iota_base + i) by (key, row id) with the uniform-hash path.  *ok = 0 when
// the input is not eligible or a bucket overflowed (nothing usable was written; the input is
// untouched).  keys_in must not alias keys_out.
int hmj_sort_uniform_try(hmj_ctx* ctx, const uint64_t* keys_in, size_t n, uint32_t iota_base,
                         uint64_t* keys_out, uint32_t* vals_out, int* ok) {
  *ok = 0;
  if (keys_in == keys_out) return HMJ_OK;
  const hmj_uniform_plan p = hmj_uniform_plan_for(n, 0);
  if (p.levels == 0 || p.qb > 16) return HMJ_OK;
  if ((reinterpret_cast<uintptr_t>(keys_in) & 15) != 0 && (p.bits[0] > 8 || p.bits[1] > 8)) return HMJ_OK;
  cudaStream_t st = ctx->stream;
  hmj_uniform_buckets B;
  HMJ_TRY(hmj_uniform_scatter(ctx, keys_in, nullptr, n, iota_base, p, 0, &B));
  bucket_offsets_kernel<<<1, 1024, 0, st>>>(B.counts, B.buckets, B.offsets);
  HMJ_LAUNCHED(ctx);
  cudaEvent_t ev;
  HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_FINISH, (double)n * 24.0, &ev));
  HMJ_TRY(hmj_launch_bucket_finish(ctx, B.keys, B.vals, B.counts, B.offsets, B.buckets, p.cap_seg,
                                   p.qb, keys_out, vals_out, B.flag));
  hmj_prof_end(ctx, ev);
  uint32_t* h = reinterpret_cast<uint32_t*>(ctx->pinned);
  HMJ_CUDA(ctx, cudaMemcpyAsync(h, B.flag, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  HMJ_CUDA(ctx, cudaStreamSynchronize(st));
  *ok = h[0] == 0;
  return HMJ_OK;
}

// ---- launch helpers of the multi-GPU join (dist_join.cu) --------------------------------
// Level 1 of the job on rows [0, n) of keys_in (a part of this rank's relation): regions are
// spread over the GPUs (peer pointers).  d_cursors: the rank's 2^bits cursors, which run on over
// the parts of a relation (zeroed by the caller before the first).  The cursors reached after this
// part are handed to the owners' count table `count_table` even when the rank has no rows (they
// must not see old counts).
int hmj_scatter_level1_dist(hmj_ctx* ctx, cudaStream_t st, const uint64_t* keys_in, size_t n,
                            uint32_t iota_base, int shift, int bits, uint32_t out_cap,
                            const hmj_peer_regions& peers, uint32_t* d_cursors, uint32_t* d_flag,
                            uint32_t* d_tile_counter, uint32_t count_table_offset,
                            uint32_t ctas_per_sm) {
  HMJ_ARG(ctx, (reinterpret_cast<uintptr_t>(keys_in) & 15) == 0 && bits <= 9);
  if (n) {
    cudaEvent_t ev;
    HMJ_TRY(hmj_prof_begin_on(ctx, HMJ_K_XSCATTER, (double)n * 20.0, &ev, st));
    ScatterLaunch a = {};
    a.keys_in = keys_in; a.n_single = (uint32_t)n; a.iota_base = iota_base;
    a.total_tiles = (uint32_t)hmj_div_up(n, tile_items_for(bits));
    a.tiles_per_region = a.total_tiles;
    a.shift = shift; a.bits = bits; a.out_cap = out_cap;
    a.cursors = d_cursors; a.flag = d_flag; a.tile_counter = d_tile_counter;
    a.peers = &peers;
    HMJ_TRY(launch_scatter(ctx, a, ctas_per_sm, st));
    hmj_prof_end_on(ctx, ev, st);
  }
  push_counts_kernel<<<((1u << bits) + 255) / 256, 256, 0, st>>>(d_cursors, 1u << bits, peers,
                                                                 count_table_offset);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

// A local level on this rank's own regions (`regions` of `cap1` slots holding the pairs
// [start1[r], cur1[r]); start1 == nullptr: from 0) -> padded buckets
// (bucket = ((region >> region_shift) << bits) | digit), appended at the buckets' cursors cur2.
int hmj_scatter_level2(hmj_ctx* ctx, cudaStream_t st, const uint64_t* k1, const uint32_t* v1,
                       const uint32_t* cur1, const uint32_t* start1, uint32_t cap1, uint32_t regions,
                       size_t n_hint, int shift, int bits, uint64_t* k2, uint32_t* v2,
                       uint32_t cap_seg, uint32_t* cur2, uint32_t* d_flag, uint32_t* d_tile_counter,
                       int region_shift) {
  cudaEvent_t ev;
  HMJ_TRY(hmj_prof_begin_on(ctx, HMJ_K_SCATTER, (double)n_hint * 24.0, &ev, st));
  ScatterLaunch a = {};
  a.keys_in = k1; a.vals_in = v1; a.in_counts = cur1; a.in_starts = start1; a.in_stride = cap1;
  // (+1 tile: a part may start up to 3 pairs before a tile boundary of the region)
  a.tiles_per_region = (uint32_t)hmj_div_up(cap1, tile_items_for(bits)) + (start1 ? 1u : 0u);
  a.total_tiles = regions * a.tiles_per_region;
  a.keys_out = k2; a.vals_out = v2;
  a.shift = shift; a.bits = bits; a.out_cap = cap_seg;
  a.cursors = cur2; a.flag = d_flag; a.tile_counter = d_tile_counter;
  a.region_shift = region_shift;
  HMJ_TRY(launch_scatter(ctx, a, 0, st));
  hmj_prof_end_on(ctx, ev, st);
  return HMJ_OK;
}
