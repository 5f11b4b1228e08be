// bucket_join.cu -- K6u: the join of two relations that were scattered into the SAME hash-prefix
// buckets by the uniform path (bucket_scatter.cu), without materialising the sorted columns.
//
// The reference sorts both relations completely and then walks the two sorted hash streams
// (hashjoin.h:56-68, 70-180).  Once both relations are cut on the same top hash bits, bucket b of
// R can only match bucket b of S, and a bucket pair (~2 x 4096 pairs) fits in shared memory.  A
// CTA therefore finishes the ordering of S's bucket on the next 12 hash bits in shared memory
// (counting sort, as segment_sort.cu), lets every R pair look at the 1-2 S pairs of its bin,
// compares the full 64-bit hashes (and, for strings, the keys: hashjoin.h:139-141) and emits
// (build_row, probe_row).  That removes the write of both sorted relations (12 B per tuple), their
// re-read by the merge join (12 B per tuple) and the merge-path search.
// Output: a bucket's pairs are staged in shared memory and written with coalesced stores at a
// position reserved with one atomicAdd on the global pair counter.  The ORDER of the pairs is
// therefore unspecified (bucket arrival order; inside a bucket the arrival order of the
// scatter); hmj_merge_join_write on sorted relations gives hash order when that is wanted.
// Algorithmic bytes: 12 per input tuple + 8 per output pair.
#include "common.cuh"
#include "join_device.cuh"

namespace {

constexpr int BJ_THREADS = 512;
constexpr int BJ_ITEMS = 9;
constexpr int BJ_CAP = BJ_THREADS * BJ_ITEMS;  // == hmj_segment_sort_cap()
constexpr int BJ_BINS = 4096;
constexpr int BJ_PER = BJ_BINS / BJ_THREADS;
constexpr size_t BJ_SMEM = (size_t)BJ_CAP * 8 + (size_t)BJ_CAP * 8 + (size_t)BJ_CAP * 4 +
                           (size_t)(BJ_BINS + 1) * 4 + 64 * 8 + 64;

struct BJArgs {
  const uint64_t* rk; const uint32_t* rv; const uint32_t* rcnt; const uint32_t* rflag;
  const uint64_t* sk; const uint32_t* sv; const uint32_t* scnt; const uint32_t* sflag;
  uint32_t in_cap, buckets;
  int dshift;              // bin of a pair = (hash >> dshift) & 4095: the 12 bits below the bucket prefix
  const uint64_t* roff; const uint8_t* rchars;  // string key storage (row-id indexed)
  const uint64_t* soff; const uint8_t* schars;
  uint32_t* out_r; uint32_t* out_s;
  uint64_t cap;            // capacity of out_r / out_s in pairs
  uint64_t* total;         // number of pairs of the whole join (zeroed; output cursor)
};

__device__ __forceinline__ uint32_t bj_smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

template <int KIND>
__global__ void __launch_bounds__(BJ_THREADS, 2)
bucket_join_kernel(const BJArgs A) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  // R hashes arrive here by a bulk async copy; once every thread holds its R pairs in registers
  // the same bytes stage the output pairs (s_or / s_os)
  uint64_t* s_rk = reinterpret_cast<uint64_t*>(smem_raw);
  uint32_t* s_or = reinterpret_cast<uint32_t*>(smem_raw);
  uint32_t* s_os = s_or + BJ_CAP;
  uint64_t* s_keys = s_rk + BJ_CAP;                           // S hashes, grouped by bin
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys + BJ_CAP);
  uint32_t* s_bin = s_vals + BJ_CAP;                          // [BINS + 1]
  uint64_t* s_tmp = reinterpret_cast<uint64_t*>(
      (reinterpret_cast<uintptr_t>(s_bin + BJ_BINS + 1) + 7) & ~uintptr_t(7));  // [64]
  uint64_t* s_bar = s_tmp + 48;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t b = blockIdx.x;

  // a scatter level overflowed somewhere: the buckets are void, the host takes the other path
  const bool dead = *reinterpret_cast<const volatile uint32_t*>(A.rflag) != 0 ||
                    *reinterpret_cast<const volatile uint32_t*>(A.sflag) != 0;
  if (dead) return;
  const uint32_t nr = min(A.rcnt[b], A.in_cap), ns = min(A.scnt[b], A.in_cap);
  if (nr == 0 || ns == 0) return;
  const uint32_t base_slot = b * A.in_cap;

  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bj_smem_u32(s_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // (the copy length is rounded up to 16 bytes: the extra slot lies inside the bucket's region)
    const uint32_t bytes = ((nr + 1u) & ~1u) * 8u;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bj_smem_u32(s_bar)),
                 "r"(bytes)
                 : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(bj_smem_u32(s_rk)),
        "l"(A.rk + base_slot), "r"(bytes), "r"(bj_smem_u32(s_bar))
        : "memory");
  }

  // ---- S bucket into registers
  uint64_t skey[BJ_ITEMS];
  uint32_t srow[BJ_ITEMS];
#pragma unroll
  for (int k = 0; k < BJ_ITEMS; ++k) {
    const uint32_t i = tid + k * BJ_THREADS;
    if (i < ns) {
      skey[k] = A.sk[base_slot + i];
      srow[k] = A.sv[base_slot + i];
    }
  }
  for (int i = tid; i <= BJ_BINS; i += BJ_THREADS) s_bin[i] = 0;
  const int s = A.dshift;
  __syncthreads();

  // ---- S: counting sort on the 12 bits below the prefix
  {
    uint16_t slot[BJ_ITEMS];
#pragma unroll
    for (int k = 0; k < BJ_ITEMS; ++k) {
      const uint32_t i = tid + k * BJ_THREADS;
      if (i < ns) slot[k] = (uint16_t)atomicAdd(&s_bin[((uint32_t)(skey[k] >> s) & (BJ_BINS - 1))], 1u);
    }
    __syncthreads();
    {
      uint32_t c[BJ_PER], sum = 0;
#pragma unroll
      for (int j = 0; j < BJ_PER; ++j) {
        c[j] = s_bin[tid * BJ_PER + j];
        sum += c[j];
      }
      uint32_t x = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      uint32_t* s_w = reinterpret_cast<uint32_t*>(s_tmp);
      if (lane == 31) s_w[warp] = x;
      __syncthreads();
      uint32_t wt = lane < warp ? s_w[lane] : 0;  // 16 warps
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) wt += __shfl_xor_sync(0xffffffffu, wt, o);
      uint32_t pre = wt + x - sum;
#pragma unroll
      for (int j = 0; j < BJ_PER; ++j) {
        s_bin[tid * BJ_PER + j] = pre;
        pre += c[j];
      }
      if (tid == 0) s_bin[BJ_BINS] = ns;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BJ_ITEMS; ++k) {
      const uint32_t i = tid + k * BJ_THREADS;
      if (i < ns) {
        const uint32_t p = s_bin[((uint32_t)(skey[k] >> s) & (BJ_BINS - 1))] + slot[k];
        s_keys[p] = skey[k];
        s_vals[p] = srow[k];
      }
    }
  }

  // ---- R pairs: hashes from the bulk copy, row ids straight from global memory (they are not
  // needed before the emit, so their latency hides behind the probe)
  {
    const uint32_t sbar = bj_smem_u32(s_bar);
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(sbar)
        : "memory");
  }
  uint64_t rkey[BJ_ITEMS];
  uint32_t rrow[BJ_ITEMS];
#pragma unroll
  for (int k = 0; k < BJ_ITEMS; ++k) {
    const uint32_t i = tid + k * BJ_THREADS;
    if (i < nr) {
      rkey[k] = s_rk[i];
      rrow[k] = A.rv[base_slot + i];
    }
  }
  __syncthreads();  // S is grouped; every thread has taken its R hashes out of the staging bytes

  // ---- probe: matches of every R pair among the S pairs of its bin.  cnt = number of matches,
  // hit = slot of the first one (the only one unless S holds duplicates of the key).
  auto match = [&](uint64_t key, uint32_t row, uint32_t j) -> bool {
    if (s_keys[j] != key) return false;
    if (KIND == HMJ_KEY_U64) return true;
    return keys_equal_str(A.roff, A.rchars, row, A.soff, A.schars, s_vals[j]);
  };
  uint32_t mine = 0;
  uint16_t cnt[BJ_ITEMS], hit[BJ_ITEMS];
#pragma unroll
  for (int k = 0; k < BJ_ITEMS; ++k) {
    const uint32_t i = tid + k * BJ_THREADS;
    cnt[k] = 0;
    hit[k] = 0;
    if (i < nr) {
      const uint32_t d = ((uint32_t)(rkey[k] >> s) & (BJ_BINS - 1));
      const uint32_t b0 = s_bin[d], n = s_bin[d + 1] - b0;
      uint32_t c = 0, first = 0;
      if (KIND == HMJ_KEY_U64) {
        // four slots at once without a branch (reads past the bin are masked), then the rest
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j) {
          const bool m = j < n && s_keys[min(b0 + j, (uint32_t)BJ_CAP - 1)] == rkey[k];
          if (m && c == 0) first = b0 + j;
          c += m ? 1u : 0u;
        }
        for (uint32_t j = 4; j < n; ++j) {
          const bool m = s_keys[b0 + j] == rkey[k];
          if (m && c == 0) first = b0 + j;
          c += m ? 1u : 0u;
        }
      } else {
        for (uint32_t j = 0; j < n; ++j) {
          const bool m = match(rkey[k], rrow[k], b0 + j);
          if (m && c == 0) first = b0 + j;
          c += m ? 1u : 0u;
        }
      }
      cnt[k] = (uint16_t)c;  // a bin holds <= CAP pairs
      hit[k] = (uint16_t)first;
      mine += c;
    }
  }
  // exclusive scan of the per-thread totals
  uint32_t t_off, tile_total;
  {
    uint32_t x = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    uint32_t* s_w = reinterpret_cast<uint32_t*>(s_tmp);
    if (lane == 31) s_w[32 + warp] = x;
    __syncthreads();
    uint32_t wt = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < BJ_THREADS / 32; ++w) {
      const uint32_t t = s_w[32 + w];
      if (w < warp) wt += t;
      tot += t;
    }
    t_off = wt + x - mine;
    tile_total = tot;
  }
  if (tile_total == 0) return;
  // ---- room for the bucket's pairs: one atomicAdd on the global pair counter (buckets land in
  // the output in the order in which they get here)
  if (tid == 0) s_tmp[40] = atomicAdd(reinterpret_cast<unsigned long long*>(A.total),
                                      (unsigned long long)tile_total);

  // ---- emit: staged in shared memory when the bucket's pairs fit, else straight to global
  const bool staged = tile_total <= (uint32_t)BJ_CAP;
  if (staged) {
    uint32_t off = t_off;
#pragma unroll
    for (int k = 0; k < BJ_ITEMS; ++k) {
      const uint32_t i = tid + k * BJ_THREADS;
      if (i < nr && cnt[k]) {
        s_or[off] = rrow[k];
        s_os[off] = s_vals[hit[k]];
        ++off;
        if (cnt[k] > 1) {
          const uint32_t d = ((uint32_t)(rkey[k] >> s) & (BJ_BINS - 1));
          const uint32_t b1 = s_bin[d + 1];
          for (uint32_t j = (uint32_t)hit[k] + 1; j < b1; ++j) {
            if (!match(rkey[k], rrow[k], j)) continue;
            s_or[off] = rrow[k];
            s_os[off] = s_vals[j];
            ++off;
          }
        }
      }
    }
  }
  __syncthreads();
  const uint64_t gbase = s_tmp[40];
  if (gbase + tile_total > A.cap) return;  // does not fit: the host re-runs with the exact size
  if (staged) {
    for (uint32_t i = tid; i < tile_total; i += BJ_THREADS) {
      A.out_r[gbase + i] = s_or[i];
      A.out_s[gbase + i] = s_os[i];
    }
  } else {
    uint64_t off = gbase + t_off;
#pragma unroll
    for (int k = 0; k < BJ_ITEMS; ++k) {
      const uint32_t i = tid + k * BJ_THREADS;
      if (i < nr && cnt[k]) {
        const uint32_t d = ((uint32_t)(rkey[k] >> s) & (BJ_BINS - 1));
        const uint32_t b1 = s_bin[d + 1];
        for (uint32_t j = hit[k]; j < b1; ++j) {
          if (!match(rkey[k], rrow[k], j)) continue;
          A.out_r[off] = rrow[k];
          A.out_s[off] = s_vals[j];
          ++off;
        }
      }
    }
  }
}


// ---- table variant ----------------------------------------------------------------------
// Same job, other data structure: instead of ordering S's bucket with a counting sort (one
// shared-memory atomic, one bin lookup and a 12-byte permutation per S pair, then two bin bounds
// and up to four 8-byte key reads per R pair), S stays where the bulk copy put it and a
// direct-mapped table of 16-bit positions is laid over it:
//   insert   every S pair writes its position into T[h(key)] with a PLAIN store (last writer wins),
//   resolve  reads the slot back; a pair that lost links itself behind the winner with one
//            atomicExch on the winner's `next` word (about a fifth of the pairs at load 0.5),
//   probe    an R pair reads T[h(key)], compares the winner's full hash (and the key, for strings)
//            and follows `next` while the chain lasts -- equal keys always share a slot, so the
//            chain holds every duplicate.
// h is a multiplicative hash of the whole 64-bit hash value, so bit patterns of identity-hashed
// integer keys do not pile up in a slot.  Shared-memory work per tuple drops to about a third and
// the S pairs never pass through registers.  First matches leave through a staging area (which
// reuses the S hashes' space) with coalesced stores; further matches of an R pair (duplicate
// keys in S) are written straight from the chain walk behind them.
constexpr uint32_t BT_NIL = 0x7FFFu;     // T entry: S position (15 bits) | 0x8000 when a chain hangs behind it
constexpr uint32_t BT_CHAIN = 0x8000u;
template <int TBITS>
struct BtCfg {
  static constexpr int TBL = 1 << TBITS;
  static constexpr size_t SMEM = (size_t)BJ_CAP * 8 + (size_t)BJ_CAP * 4 + (size_t)BJ_CAP * 4 +
                                 (size_t)TBL * 2 + 64 * 8 + 64;
  static_assert(2 * (SMEM + 1024) <= 233472, "two CTAs of the table join per SM");
};

template <int TBITS>
__device__ __forceinline__ uint32_t bt_slot(uint64_t key) {
  return (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> (64 - TBITS));
}

template <int KIND, int TBITS>
__global__ void __launch_bounds__(BJ_THREADS, 2)
bucket_join_table_kernel(const BJArgs A) {
  constexpr int TBL = BtCfg<TBITS>::TBL;
  extern __shared__ __align__(128) unsigned char smem_bt[];
  uint64_t* s_sk = reinterpret_cast<uint64_t*>(smem_bt);            // S hashes as they arrived
  uint32_t* s_or = reinterpret_cast<uint32_t*>(smem_bt);            // ... later the staged pairs
  uint32_t* s_os = s_or + BJ_CAP;
  uint32_t* s_sv = reinterpret_cast<uint32_t*>(s_sk + BJ_CAP);      // S row ids
  uint32_t* s_nxt = s_sv + BJ_CAP;                                  // chain: next S position
  uint16_t* s_T = reinterpret_cast<uint16_t*>(s_nxt + BJ_CAP);      // slot -> S position
  uint64_t* s_tmp = reinterpret_cast<uint64_t*>(s_T + TBL);         // [64]
  uint64_t* s_bar = s_tmp + 48;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t b = blockIdx.x;

  // (flags and counts are independent loads: one round trip, not two)
  const uint32_t rflag = *reinterpret_cast<const volatile uint32_t*>(A.rflag);
  const uint32_t sflag = *reinterpret_cast<const volatile uint32_t*>(A.sflag);
  const uint32_t nr = min(A.rcnt[b], A.in_cap), ns = min(A.scnt[b], A.in_cap);
  if ((rflag | sflag) != 0) return;  // a scatter level overflowed: the buckets are void
  if (nr == 0 || ns == 0) return;
  const uint32_t base_slot = b * A.in_cap;

  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bj_smem_u32(s_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // (lengths rounded up to 16 bytes: the extra slots lie inside the bucket's region)
    const uint32_t kbytes = ((ns + 1u) & ~1u) * 8u, vbytes = ((ns + 3u) & ~3u) * 4u;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bj_smem_u32(s_bar)),
                 "r"(kbytes + vbytes)
                 : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(bj_smem_u32(s_sk)),
        "l"(A.sk + base_slot), "r"(kbytes), "r"(bj_smem_u32(s_bar))
        : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(bj_smem_u32(s_sv)),
        "l"(A.sv + base_slot), "r"(vbytes), "r"(bj_smem_u32(s_bar))
        : "memory");
  }
  // R hashes straight into registers: in flight while the table is built
  uint64_t rkey[BJ_ITEMS];
  uint32_t rrow[BJ_ITEMS];
#pragma unroll
  for (int k = 0; k < BJ_ITEMS; ++k) {
    const uint32_t i = tid + k * BJ_THREADS;
    rkey[k] = i < nr ? __ldcs(A.rk + base_slot + i) : 0;
    if (KIND != HMJ_KEY_U64) rrow[k] = i < nr ? __ldcs(A.rv + base_slot + i) : 0;
  }
  {
    uint4* t4 = reinterpret_cast<uint4*>(s_T);
    const uint32_t nil2 = BT_NIL | (BT_NIL << 16);
    const uint4 nil = make_uint4(nil2, nil2, nil2, nil2);
    for (int i = tid; i < TBL * 2 / 16; i += BJ_THREADS) t4[i] = nil;
    uint4* n4 = reinterpret_cast<uint4*>(s_nxt);
    for (int i = tid; i < (int)((ns + 3u) / 4u); i += BJ_THREADS) n4[i] = nil;  // S positions only; low half = NIL is what counts
  }
  __syncthreads();
  {
    const uint32_t sbar = bj_smem_u32(s_bar);
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(sbar)
        : "memory");
  }

  // ---- insert: last writer wins
  uint16_t slot_of[BJ_ITEMS];
#pragma unroll
  for (int k = 0; k < BJ_ITEMS; ++k) {
    const uint32_t i = tid + k * BJ_THREADS;
    if (i < ns) {
      slot_of[k] = (uint16_t)bt_slot<TBITS>(s_sk[i]);
      s_T[slot_of[k]] = (uint16_t)i;
    }
  }
  __syncthreads();
  // ---- resolve: pairs that lost their slot chain up behind the winner and mark the slot
  // (every loser stores the same marked value: a benign race)
#pragma unroll
  for (int k = 0; k < BJ_ITEMS; ++k) {
    const uint32_t i = tid + k * BJ_THREADS;
    if (i < ns) {
      const uint32_t w = s_T[slot_of[k]] & BT_NIL;
      if (w != i) {
        s_nxt[i] = atomicExch(&s_nxt[w], i);
        s_T[slot_of[k]] = (uint16_t)(w | BT_CHAIN);
      }
    }
  }
  __syncthreads();

  // ---- probe: cnt = matches of the R pair, hit = S position of the first one
  auto match = [&](uint64_t key, uint32_t row, uint32_t j) -> bool {
    if (s_sk[j] != key) return false;
    if (KIND == HMJ_KEY_U64) return true;
    return keys_equal_str(A.roff, A.rchars, row, A.soff, A.schars, s_sv[j]);
  };
  uint16_t cnt[BJ_ITEMS], hit[BJ_ITEMS];
  uint32_t firsts = 0, extras = 0;
#pragma unroll
  for (int k = 0; k < BJ_ITEMS; ++k) {
    const uint32_t i = tid + k * BJ_THREADS;
    uint32_t c = 0, first = BT_NIL;
    if (i < nr) {
      const uint32_t e = s_T[bt_slot<TBITS>(rkey[k])];
      uint32_t j = e & BT_NIL;
      if (j != BT_NIL) {
        if (match(rkey[k], KIND != HMJ_KEY_U64 ? rrow[k] : 0u, j)) {
          first = j;
          c = 1;
        }
        if (e & BT_CHAIN) {
          j = s_nxt[j] & BT_NIL;
          while (j != BT_NIL) {
            if (match(rkey[k], KIND != HMJ_KEY_U64 ? rrow[k] : 0u, j)) {
              if (c == 0) first = j;
              ++c;
            }
            j = s_nxt[j] & BT_NIL;
          }
        }
      }
    }
    cnt[k] = (uint16_t)c;  // a chain holds <= CAP pairs
    hit[k] = (uint16_t)first;
    firsts += c ? 1u : 0u;
    extras += c ? c - 1u : 0u;
  }
  // the row ids of R are not needed before the emit: their latency hides behind the probe
  if (KIND == HMJ_KEY_U64) {
#pragma unroll
    for (int k = 0; k < BJ_ITEMS; ++k) {
      const uint32_t i = tid + k * BJ_THREADS;
      rrow[k] = (i < nr && cnt[k]) ? __ldcs(A.rv + base_slot + i) : 0;
    }
  }
  // exclusive scans of the per-thread first-match and further-match counts (one 64-bit scan)
  uint32_t f_off, e_off, F, E;
  {
    const uint64_t mine = ((uint64_t)extras << 32) | firsts;
    uint64_t x = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_tmp[warp] = x;
    __syncthreads();
    uint64_t wt = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < BJ_THREADS / 32; ++w) {
      const uint64_t t = s_tmp[w];
      if (w < warp) wt += t;
      tot += t;
    }
    const uint64_t ex = wt + x - mine;
    f_off = (uint32_t)ex;
    e_off = (uint32_t)(ex >> 32);
    F = (uint32_t)tot;
    E = (uint32_t)(tot >> 32);
  }
  const uint32_t tile_total = F + E;
  if (tile_total == 0) return;
  // room for the bucket's pairs: one atomicAdd on the global pair counter.  Its round trip is
  // needed first by the stores of the further matches; without any (unique keys in S) it hides
  // behind the staging of the first matches and is read at the barrier that ends the staging.
  if (tid == 0) s_tmp[40] = atomicAdd(reinterpret_cast<unsigned long long*>(A.total),
                                      (unsigned long long)tile_total);
  // ---- further matches (duplicate keys in S): straight from the chain walk, behind the firsts
  if (E) {
    __syncthreads();
    const uint64_t gbase_e = s_tmp[40];
    if (gbase_e + tile_total > A.cap) return;  // does not fit: the host re-runs with the exact size
    uint64_t off = gbase_e + F + e_off;
#pragma unroll
    for (int k = 0; k < BJ_ITEMS; ++k) {
      if (cnt[k] > 1) {
        uint32_t j = s_nxt[hit[k]] & BT_NIL;
        while (j != BT_NIL) {
          if (match(rkey[k], rrow[k], j)) {
            A.out_r[off] = rrow[k];
            A.out_s[off] = s_sv[j];
            ++off;
          }
          j = s_nxt[j] & BT_NIL;
        }
      }
    }
    __syncthreads();  // the S hashes are dead from here on
  }
  // ---- first matches: staged, then coalesced stores.  A thread's pairs take consecutive staging
  // slots and most threads hold exactly 8 of them (4096 R pairs over 512 threads), so the lanes
  // of a warp would store with a stride of 8 words -- an 8-way bank conflict that was 27 % of the
  // kernel's shared-memory wavefronts (profiles/r2s_join_table_digest.txt).  XOR-ing the low five
  // index bits with the next five spreads such strides over all banks and keeps consecutive
  // slots (the read-back below) conflict-free.
  auto swz = [](uint32_t i) -> uint32_t { return i ^ ((i >> 5) & 31u); };
  {
    uint32_t off = f_off;
#pragma unroll
    for (int k = 0; k < BJ_ITEMS; ++k) {
      if (cnt[k]) {
        const uint32_t sv = s_sv[hit[k]];
        s_or[swz(off)] = rrow[k];
        s_os[swz(off)] = sv;
        ++off;
      }
    }
  }
  __syncthreads();
  const uint64_t gbase = s_tmp[40];
  if (gbase + tile_total > A.cap) return;  // does not fit: the host re-runs with the exact size
  for (uint32_t i = tid; i < F; i += BJ_THREADS) {
    A.out_r[gbase + i] = s_or[swz(i)];
    A.out_s[gbase + i] = s_os[swz(i)];
  }
}

}  // namespace

// Enqueues the join of the padded buckets of two relations (same bucket numbering) on the
// context's stream: pairs are appended to out_r / out_s (capacity `cap` pairs) at the position
// the device counter *d_total holds, which the kernel advances.  Nothing is zeroed or read back.
int hmj_bucket_join_launch(hmj_ctx* ctx, int kind, const hmj_uniform_buckets& R,
                           const hmj_uniform_buckets& S, uint32_t cap_seg, int dshift, size_t nr,
                           size_t ns, const uint64_t* roff, const uint8_t* rchars,
                           const uint64_t* soff, const uint8_t* schars, uint32_t* out_r,
                           uint32_t* out_s, size_t cap, uint64_t* d_total) {
  cudaStream_t st = ctx->stream;
  BJArgs A = {};
  A.rk = R.keys; A.rv = R.vals; A.rcnt = R.counts; A.rflag = R.flag;
  A.sk = S.keys; A.sv = S.vals; A.scnt = S.counts; A.sflag = S.flag;
  A.in_cap = cap_seg;
  A.buckets = R.buckets;
  A.dshift = dshift;
  A.roff = roff; A.rchars = rchars; A.soff = soff; A.schars = schars;
  A.total = d_total;
  A.out_r = out_r; A.out_s = out_s; A.cap = cap;
  cudaEvent_t ev;
  HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_BJOIN, ((double)nr + (double)ns) * 12.0, &ev));
  // HMJ_BUCKET_JOIN=sort selects the counting-sort variant (kept for comparison)
  static const bool table = [] {
    const char* e = getenv("HMJ_BUCKET_JOIN");
    return !(e && strcmp(e, "sort") == 0);
  }();
  static const int tbits = [] {
    const char* e = getenv("HMJ_BUCKET_TABLE_BITS");
    return e && atoi(e) == 14 ? 14 : 13;  // measured on C3: 2.67 ms with 2^13 slots, 2.70 with 2^14
  }();
#define HMJ_BT_LAUNCH(K, TB)                                                                        \
  do {                                                                                              \
    HMJ_CUDA(ctx, cudaFuncSetAttribute(bucket_join_table_kernel<K, TB>,                             \
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,                 \
                                       (int)BtCfg<TB>::SMEM));                                      \
    bucket_join_table_kernel<K, TB><<<R.buckets, BJ_THREADS, BtCfg<TB>::SMEM, st>>>(A);             \
  } while (0)
  if (table && kind == HMJ_KEY_U64) {
    if (tbits == 13) HMJ_BT_LAUNCH(HMJ_KEY_U64, 13); else HMJ_BT_LAUNCH(HMJ_KEY_U64, 14);
  } else if (table) {
    if (tbits == 13) HMJ_BT_LAUNCH(HMJ_KEY_STRING, 13); else HMJ_BT_LAUNCH(HMJ_KEY_STRING, 14);
  } else if (kind == HMJ_KEY_U64) {
    HMJ_CUDA(ctx, cudaFuncSetAttribute(bucket_join_kernel<HMJ_KEY_U64>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BJ_SMEM));
    bucket_join_kernel<HMJ_KEY_U64><<<R.buckets, BJ_THREADS, BJ_SMEM, st>>>(A);
  } else {
    HMJ_CUDA(ctx, cudaFuncSetAttribute(bucket_join_kernel<HMJ_KEY_STRING>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BJ_SMEM));
    bucket_join_kernel<HMJ_KEY_STRING><<<R.buckets, BJ_THREADS, BJ_SMEM, st>>>(A);
  }
#undef HMJ_BT_LAUNCH
  HMJ_LAUNCHED(ctx);
  hmj_prof_end(ctx, ev);
  return HMJ_OK;
}

// Joins the padded buckets of two relations (same bucket numbering).  *ok = 0 when a scatter level
// overflowed; on success the pairs are in WS_OUT_R / WS_OUT_S and *h_num_pairs is their number.
int hmj_bucket_join_run(hmj_ctx* ctx, int kind, const hmj_uniform_buckets& R,
                        const hmj_uniform_buckets& S, uint32_t cap_seg, int dshift, size_t nr,
                        size_t ns, const uint64_t* roff, const uint8_t* rchars,
                        const uint64_t* soff, const uint8_t* schars, uint64_t* h_num_pairs,
                        int* ok) {
  *ok = 0;
  cudaStream_t st = ctx->stream;
  // the global pair counter
  uint64_t* d_total = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_JOIN_TILES, 4, &d_total));

  // optimistic capacity: what the buffers already hold, at least max(|R|,|S|) pairs
  size_t cap = ctx->ws[WS_OUT_R].bytes / 4;
  const size_t guess = nr > ns ? nr : ns;
  if (cap < guess) cap = guess;
  uint32_t *out_r = nullptr, *out_s = nullptr;
  uint64_t* h = reinterpret_cast<uint64_t*>(ctx->pinned);
  for (int attempt = 0; attempt < 2; ++attempt) {
    HMJ_TRY(hmj_ws(ctx, WS_OUT_R, cap, &out_r));
    HMJ_TRY(hmj_ws(ctx, WS_OUT_S, cap, &out_s));
    cap = ctx->ws[WS_OUT_R].bytes / 4;
    if (ctx->ws[WS_OUT_S].bytes / 4 < cap) cap = ctx->ws[WS_OUT_S].bytes / 4;
    HMJ_TRY(hmj_fill_async(ctx, d_total, 0, 32, st));
    HMJ_TRY(hmj_bucket_join_launch(ctx, kind, R, S, cap_seg, dshift, nr, ns, roff, rchars, soff,
                                   schars, out_r, out_s, cap, d_total));
    // [0] total pairs, [1] R flag | S flag << 32
    HMJ_CUDA(ctx, cudaMemcpyAsync(h, d_total, 8, cudaMemcpyDeviceToHost, st));
    HMJ_CUDA(ctx, cudaMemcpyAsync(reinterpret_cast<uint32_t*>(h + 1), R.flag, 4, cudaMemcpyDeviceToHost, st));
    HMJ_CUDA(ctx, cudaMemcpyAsync(reinterpret_cast<uint32_t*>(h + 1) + 1, S.flag, 4, cudaMemcpyDeviceToHost, st));
    HMJ_CUDA(ctx, cudaStreamSynchronize(st));
    if (h[1] != 0) return HMJ_OK;  // overflow: not ok
    const uint64_t m = h[0];
    if (m <= cap) {
      if (ctx->prof_on && ctx->prof_used >= 2) ctx->prof_bytes[ctx->prof_used / 2 - 1] += 8.0 * (double)m;
      ctx->last_pairs = m;
      *h_num_pairs = m;
      *ok = 1;
      return HMJ_OK;
    }
    cap = (size_t)m;  // the buckets are still valid: join again into larger buffers
  }
  return hmj_fail(ctx, HMJ_ECAPACITY, "bucket join: result did not fit twice", __FILE__, __LINE__);
}

namespace {
constexpr uint32_t SAMPLE_PER_SIDE = 32768;

__global__ void __launch_bounds__(256)
sample_keys_kernel(const uint64_t* __restrict__ r, size_t nr, const uint64_t* __restrict__ s, size_t ns,
                   uint64_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < SAMPLE_PER_SIDE) {
    out[i] = r[(size_t)((unsigned __int128)i * nr / SAMPLE_PER_SIDE)];
  } else if (i < 2 * SAMPLE_PER_SIDE) {
    out[i] = s[(size_t)((unsigned __int128)(i - SAMPLE_PER_SIDE) * ns / SAMPLE_PER_SIDE)];
  }
}
}  // namespace

// The plan a sample asks for (host only; tests/test_plan_from_sample.py drives it without a GPU
// through hmj_debug_plan_from_sample).  h = per_side evenly spaced hashes of R, then of S.
//  * prefix: the top bits all sampled hashes share are skipped (and checked by level 1);
//  * density: when the hashes under the prefix fill only a part [lo, hi] of the range -- dense
//    ids 0..N-1 with N not a power of two, identity-hashed -- a region of the filled part holds
//    range / (hi - lo + 1) times its even share.  The plan then provides for n * density rows
//    (bucket count, level-1 regions).  Only for at most one row per key value (the 2^12 values of
//    a final bucket must fit its slots) and density <= 4;
//  * skew: a level-1 region with more than its share (x density) + 50 % + 64 of a side's sample.
void hmj_plan_from_sample(const uint64_t* h, uint32_t per_side, size_t nr, size_t ns,
                          hmj_uniform_plan* p, bool skewed_side[2]) {
  skewed_side[0] = skewed_side[1] = false;
  const size_t n_side[2] = {nr, ns};
  uint64_t diff = 0, lo[2] = {~0ull, ~0ull}, hi[2] = {0, 0};
  for (uint32_t i = 0; i < 2 * per_side; ++i) diff |= h[i] ^ h[0];
  int skip = diff ? __builtin_clzll(diff) : 64;
  const size_t n_plan = nr > ns ? nr : ns;
  if (hmj_uniform_plan_for(n_plan, 0).levels == 0) return;  // (*p stays the caller's)
  // how much of the range under the prefix the keys fill, per side
  const long double range = ldexpl(1.0L, 64 - skip);
  long double dens[2] = {1.0L, 1.0L}, density = 1.0L, rows_per_value = 0.0L;
  // (only under a shared prefix: hashes that use all 64 bits are not looked at again -- this runs
  // inside every sampled join)
  for (int side = 0; side < 2 && skip > 0; ++side) {
    const uint64_t* hs = h + (size_t)side * per_side;
    uint64_t l = ~0ull, m = 0;
    for (uint32_t i = 0; i < per_side; ++i) {
      const uint64_t v = hs[i];
      l = v < l ? v : l;
      m = v > m ? v : m;
    }
    lo[side] = l;
    hi[side] = m;
    const long double span = (long double)(hi[side] - lo[side]) + 1.0L;
    dens[side] = range / span;
    if (dens[side] > density) density = dens[side];
    if ((long double)n_side[side] / span > rows_per_value) rows_per_value = (long double)n_side[side] / span;
  }
  size_t n_eff = n_plan;
  if (density > 1.001L && density <= 4.0L && rows_per_value <= 1.001L &&
      (long double)n_plan * density < 2147483648.0L) {
    long double want = (long double)n_plan * density;
    if (want > range) want = range;  // (<= one row per value: the sample's span is a hair short)
    n_eff = (size_t)ceill(want);
  } else {
    density = 1.0L;
    dens[0] = dens[1] = 1.0L;
  }
  hmj_uniform_plan q = hmj_uniform_plan_for(n_eff, 0);
  if (q.levels == 0) return;
  if (skip > 64 - q.qb - 12) skip = 64 - q.qb - 12;
  if (skip > 0) {
    q = hmj_uniform_plan_for(n_eff, skip);
    if (q.levels == 0) {
      *p = q;
      return;
    }
    q.check_prefix = true;
    q.prefix = h[0] >> (64 - skip);
  }
  q.density = (double)density;
  // share of the sample per level-1 region
  uint32_t hist[2][512] = {};
  const int shift = 64 - q.skip - q.bits[0];
  const uint32_t mask = (1u << q.bits[0]) - 1;
  for (int side = 0; side < 2; ++side) {
    const uint64_t* hs = h + (size_t)side * per_side;
    uint32_t* hh = hist[side];
    for (uint32_t i = 0; i < per_side; ++i) hh[(uint32_t)(hs[i] >> shift) & mask]++;
  }
  for (int side = 0; side < 2; ++side) {
    const uint32_t expect = (uint32_t)((long double)(per_side >> q.bits[0]) * dens[side]);
    for (uint32_t b = 0; b <= mask; ++b)
      if (hist[side][b] > expect + expect / 2 + 64) skewed_side[side] = true;
  }
  *p = q;
}

extern "C" int hmj_debug_plan_from_sample(const uint64_t* h_sample, uint32_t per_side, size_t nr,
                                          size_t ns, int64_t* out10) {
  hmj_uniform_plan p = hmj_uniform_plan_for(nr > ns ? nr : ns, 0);
  bool sk[2];
  hmj_plan_from_sample(h_sample, per_side, nr, ns, &p, sk);
  out10[0] = p.levels; out10[1] = p.bits[0]; out10[2] = p.bits[1]; out10[3] = p.qb; out10[4] = p.skip;
  out10[5] = p.cap_seg; out10[6] = p.check_prefix; out10[7] = (int64_t)p.prefix;
  out10[8] = (int64_t)llround((p.density > 1.0 ? p.density : 1.0) * 1e6);
  out10[9] = (sk[0] ? 1 : 0) | (sk[1] ? 2 : 0);
  return 0;
}

// Looks at 32768 evenly spaced hashes of each relation before anything is moved:
//  * top bits that every sampled hash of BOTH relations shares (identity-hashed dense integer keys
//    0..N-1 share their upper 64 - log2(N) bits) are skipped by the bucket plan, so such keys cut
//    as evenly as random ones; level 1 verifies the assumption on every row;
//  * a level-1 region that holds more than 1.5x its share of the sample (Zipf heads, heavy
//    duplicates) means the padded buckets would overflow: *skewed = 1 and the caller goes to the
//    sort + merge-path plan without paying for a failed scatter attempt.
// Costs one small kernel, a 512 KB copy and a stream synchronisation (~0.1 ms), so only joins of
// at least 2^22 rows are sampled.
int hmj_sample_plan(hmj_ctx* ctx, const uint64_t* r_hash, size_t nr, const uint64_t* s_hash,
                    size_t ns, hmj_uniform_plan* p, int* skewed) {
  *skewed = 0;
  cudaStream_t st = ctx->stream;
  uint64_t* d_sample = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_JOIN_MISC, 2 * SAMPLE_PER_SIDE, &d_sample));
  if (ctx->sample_host.size() < 2 * SAMPLE_PER_SIDE) {
    // (pinned once per context; the join never allocates in steady state)
    ctx->sample_host.resize(2 * SAMPLE_PER_SIDE);
    cudaHostRegister(ctx->sample_host.data(), ctx->sample_host.size() * 8, cudaHostRegisterDefault);
    cudaGetLastError();
  }
  sample_keys_kernel<<<2 * SAMPLE_PER_SIDE / 256, 256, 0, st>>>(r_hash, nr, s_hash, ns, d_sample);
  HMJ_LAUNCHED(ctx);
  uint64_t* h = ctx->sample_host.data();
  HMJ_CUDA(ctx, cudaMemcpyAsync(h, d_sample, 2 * SAMPLE_PER_SIDE * 8, cudaMemcpyDeviceToHost, st));
  HMJ_CUDA(ctx, cudaStreamSynchronize(st));
  bool sk[2] = {false, false};
  hmj_plan_from_sample(h, SAMPLE_PER_SIDE, nr, ns, p, sk);
  for (int side = 0; side < 2; ++side) {
    ctx->sample_skewed[side] = sk[side];
    if (sk[side]) *skewed = 1;
  }
  return HMJ_OK;
}

// Whole-operator join through the uniform path: scatter R and S into the same 2^qb buckets, join
// bucket by bucket.  *ok = 0 when the inputs are not eligible or a bucket overflowed (the host
// then sorts both relations and merge-joins them); on success the pairs are in the context's
// result buffers (WS_OUT_R / WS_OUT_S) and *h_num_pairs is their number.
// kind = HMJ_KEY_U64: r_hash / s_hash are the key columns themselves.  HMJ_KEY_STRING: they are
// the hash columns and the offsets/chars arrays give the keys for verification.
// skip_bits > 0: the caller knows that all hashes share that many top bits (a rank's hash range).
int hmj_join_uniform_try(hmj_ctx* ctx, int kind, const uint64_t* r_hash, const uint32_t* r_rowid,
                         size_t nr, const uint64_t* s_hash, const uint32_t* s_rowid, size_t ns,
                         int skip_bits, const uint64_t* roff,
                         const uint8_t* rchars, const uint64_t* soff, const uint8_t* schars,
                         cudaEvent_t ev_partitioned, uint64_t* h_num_pairs, int* ok) {
  *ok = 0;
  ctx->sample_skewed[0] = ctx->sample_skewed[1] = false;
  if (nr == 0 || ns == 0) return HMJ_OK;
  hmj_uniform_plan p = hmj_uniform_plan_for(nr > ns ? nr : ns, skip_bits);
  if (p.levels == 0) return HMJ_OK;
  const char* e = getenv("HMJ_JOIN_UNIFORM");
  if (e && atoi(e) == 0) return HMJ_OK;
  if (ctx->hash_order) return HMJ_OK;  // the caller wants hash order: sort + merge-path plan
  static const bool sampling = [] {
    const char* s = getenv("HMJ_JOIN_SAMPLE");
    return !(s && atoi(s) == 0);
  }();
  if (sampling && skip_bits == 0 && nr + ns >= ((size_t)1 << 22) && nr >= SAMPLE_PER_SIDE && ns >= SAMPLE_PER_SIDE &&
      ((reinterpret_cast<uintptr_t>(r_hash) | reinterpret_cast<uintptr_t>(s_hash)) & 15) == 0) {
    int skewed = 0;
    HMJ_TRY(hmj_sample_plan(ctx, r_hash, nr, s_hash, ns, &p, &skewed));
    if (skewed || p.levels == 0) return HMJ_OK;
  }
  if ((p.bits[0] > 8 || p.bits[1] > 8) &&
      ((reinterpret_cast<uintptr_t>(r_hash) | reinterpret_cast<uintptr_t>(s_hash) |
        reinterpret_cast<uintptr_t>(r_rowid) | reinterpret_cast<uintptr_t>(s_rowid)) & 15) != 0)
    return HMJ_OK;
  hmj_uniform_buckets R, S;
  HMJ_TRY(hmj_uniform_scatter(ctx, r_hash, r_rowid, nr, 0, p, 0, &R));
  HMJ_TRY(hmj_uniform_scatter(ctx, s_hash, s_rowid, ns, 0, p, 1, &S));
  if (ev_partitioned) HMJ_CUDA(ctx, cudaEventRecord(ev_partitioned, ctx->stream));
  return hmj_bucket_join_run(ctx, kind, R, S, p.cap_seg, 64 - p.skip - p.qb - 12, nr, ns, roff,
                             rchars, soff, schars, h_num_pairs, ok);
}
