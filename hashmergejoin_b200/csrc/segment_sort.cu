// segment_sort.cu -- K3b: finishes a prefix-ordered array with ONE more read + write of the pairs.
//
// After the stable scatter passes over the top digits (radix_sort.cu) the array is ordered on
// prefix = key >> (64 - q).  The reference reaches the same state after its top-level scatter
// (radix_hash.h:351-423) and then recurses into every bucket with further histogram + permute
// levels and an insertion sort for small buckets (radix_hash.h:122-292,86-120).  Here a CTA
// takes a WINDOW of whole prefix segments (<= CAP pairs), sorts it completely in shared memory
// and writes it back in place, so the remaining 64 - q key bits cost no further HBM passes:
//   1. counting sort on a 12-bit order-preserving digit d = (key - window_base) >> s: one
//      shared-memory atomicAdd per pair gives (bin, slot-in-bin) -- no warp ranking is needed
//      because the order inside a bin is restored in step 3;
//   2. exclusive scan of the 4096 bin counts;
//   3. every pair ranks itself among its bin-mates on (full key, position in the window); hashes
//      are uniform, so a bin holds ~1-2 pairs.  Ties on the full key keep their window order,
//      which keeps the whole sort stable;
//   4. pairs go to their final slot in shared memory and leave with coalesced stores.
// Windows come from segment_windows_kernel: mode A = one window per prefix value (segments of
// ~2^11..2^12 pairs), mode B = windows of W consecutive pairs extended to the next segment
// boundary (small segments).  A window longer than CAP, or a bin with more than BIN_LIMIT
// pairs, raises *flag and leaves the window untouched; the host then re-sorts with scatter
// passes over every digit (same recovery as the bucket finish).
// Algorithmic bytes per pair: 12 read + 12 written.
#include "common.cuh"

namespace {

// 8-byte shared-memory load executed only where `p` holds; elsewhere the result is all ones
__device__ __forceinline__ uint64_t lds64_if(bool p, const uint64_t* a) {
  uint64_t v = ~0ull;
  asm volatile(
      "{\n\t"
      ".reg .pred q;\n\t"
      "setp.ne.u32 q, %2, 0;\n\t"
      "@q ld.shared.u64 %0, [%1];\n\t"
      "}"
      : "+l"(v)
      : "r"((uint32_t)__cvta_generic_to_shared(a)), "r"((uint32_t)p)
      : "memory");
  return v;
}

constexpr int SEG_BINS = 4096;
constexpr int SEG_BIN_LIMIT = 192;

template <int THREADS, int ITEMS>
struct SegCfg {
  static constexpr int CAP = THREADS * ITEMS;
  static constexpr size_t SMEM_BYTES =
      (size_t)(CAP + 4) * 8 + (size_t)CAP * 4 + (size_t)(SEG_BINS + 1) * 4 + 32 * 4 + 16;
};

// PADDED = false: window w = [win[w], win[w+1]) of (keys, vals), sorted in place; ties keep their
//   window order (stable).
// PADDED = true : window w = bucket w of the uniform sort (bucket_scatter.cu): counts[w] pairs at
//   keys[w * in_cap ...], all sharing the q-bit prefix w; written to out_* at offsets[w]; ties
//   are ordered by the payload (the row id).
template <int THREADS, int ITEMS, int MIN_BLOCKS, bool PADDED>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
segment_sort_kernel(uint64_t* __restrict__ keys, uint32_t* __restrict__ vals,
                    const uint32_t* __restrict__ win, int q, uint32_t* __restrict__ flag,
                    const uint32_t* __restrict__ counts, uint32_t in_cap,
                    uint64_t* __restrict__ out_keys, uint32_t* __restrict__ out_vals) {
  using Cfg = SegCfg<THREADS, ITEMS>;
  constexpr int CAP = Cfg::CAP;
  constexpr int PER = (SEG_BINS + THREADS - 1) / THREADS;  // bins scanned per thread
  constexpr int IDX_BITS = 13;
  static_assert(CAP <= (1 << IDX_BITS), "window positions must fit IDX_BITS");

  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem_raw);
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys + CAP + 4);
  uint16_t* s_idx = reinterpret_cast<uint16_t*>(s_vals);  // dead before s_vals is written
  uint32_t* s_bin = s_vals + CAP;                         // [SEG_BINS + 1] counts, then starts
  uint32_t* s_warp = s_bin + SEG_BINS + 1;                // [32]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t start, len, out_start;
  if (PADDED) {
    // a scatter level overflowed: counts and offsets mean nothing, the host re-sorts
    if (*reinterpret_cast<volatile uint32_t*>(flag) != 0) return;
    len = counts[blockIdx.x];
    start = blockIdx.x * in_cap;
    out_start = win[blockIdx.x];
    if (len == 0) return;
    if (len > in_cap) return;  // overflowed bucket: the scatter has raised the flag already
  } else {
    start = win[blockIdx.x];
    const uint32_t end = win[blockIdx.x + 1];
    if (end <= start + 1) return;  // empty or a single pair: already in place
    len = end - start;
    out_start = start;
    out_keys = keys;
    out_vals = vals;
  }
  if (len > (uint32_t)CAP) {
    if (tid == 0) atomicOr(flag, 1u);
    return;
  }

  for (int i = tid; i <= SEG_BINS; i += THREADS) s_bin[i] = 0;

  uint64_t key[ITEMS];
  uint32_t val[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const uint32_t i = tid + k * THREADS;
    if (i < len) {
      key[k] = keys[start + i];
      val[k] = vals[start + i];
    }
  }
  // order-preserving digit over the key range of the window's prefixes
  const int pshift = 64 - q;  // q in [0, 56]
  uint64_t base = 0;
  int s;
  if (PADDED) {
    s = max(0, pshift - 12);
    base = q ? ((uint64_t)blockIdx.x << pshift) : 0;
  } else {
    const uint64_t k_first = keys[start], k_last = keys[start + len - 1];
    const uint64_t plo = q ? (k_first >> pshift) : 0, phi = q ? (k_last >> pshift) : 0;
    const uint64_t npre = phi - plo + 1;
    if (npre <= (uint64_t)SEG_BINS) {
      const int kb = 31 - __clz((uint32_t)(SEG_BINS / (uint32_t)npre));
      s = pshift - kb;
    } else {
      const uint64_t x = (npre - 1) >> 12;
      s = pshift + (64 - __clzll((long long)x));
    }
    s = max(0, min(s, 63));
    base = q ? (plo << pshift) : 0;
  }
  __syncthreads();

  uint16_t slot[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const uint32_t i = tid + k * THREADS;
    if (i < len) {
      const uint32_t d = (uint32_t)((key[k] - base) >> s);
      slot[k] = (uint16_t)atomicAdd(&s_bin[d], 1u);
    }
  }
  __syncthreads();

  // exclusive scan of the bin counts (in place); a crowded bin sends the sort to the fallback
  {
    uint32_t c[PER], sum = 0, mx = 0;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int b = tid * PER + j;
      c[j] = b < SEG_BINS ? s_bin[b] : 0;
      sum += c[j];
      mx = max(mx, c[j]);
    }
    uint32_t x = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    const int crowded = __syncthreads_or(mx > (uint32_t)SEG_BIN_LIMIT);
    if (crowded) {
      if (tid == 0) atomicOr(flag, 1u);
      return;
    }
    // totals of the warps before mine, summed with shuffles (THREADS / 32 <= 32 warps)
    uint32_t wt = (lane < warp && lane < THREADS / 32) ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wt += __shfl_xor_sync(0xffffffffu, wt, o);
    uint32_t pre = wt + x - sum;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int b = tid * PER + j;
      if (b < SEG_BINS) s_bin[b] = pre;
      pre += c[j];
    }
    if (tid == 0) s_bin[SEG_BINS] = len;
  }
  __syncthreads();

  if (s <= 64 - IDX_BITS) {
    // fast path: one word per pair = (key bits below the digit) << IDX_BITS | window position.
    // It is unique and orders bin-mates exactly like (key, position), so ranking is one
    // comparison per bin-mate.
    const uint64_t rem_mask = (1ull << s) - 1;
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const uint32_t i = tid + k * THREADS;
      if (i < len) {
        const uint64_t rel = key[k] - base;
        const uint32_t d = (uint32_t)(rel >> s);
        s_keys[s_bin[d] + slot[k]] = ((rel & rem_mask) << IDX_BITS) | i;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const uint32_t i = tid + k * THREADS;
      if (i < len) {
        const uint64_t rel = key[k] - base;
        const uint32_t d = (uint32_t)(rel >> s);
        const uint64_t me = ((rel & rem_mask) << IDX_BITS) | i;
        const uint32_t b0 = s_bin[d], c = s_bin[d + 1] - b0;
        // The first four bin-mates without a branch, but each load PREDICATED on "this slot holds
        // another pair of my bin": a bin has one pair on average, so three of four loads drop
        // out and with them their share of the random-access bank conflicts (the unconditional
        // form of this loop was 58 M of the kernel's 141 M shared-memory wavefronts at 64M pairs,
        // profiles/r2t_segment_sort_*).  An unloaded word stays all ones, which no `me` is.
        const uint32_t own = slot[k];
        uint32_t r = 0;
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j) {
          const uint64_t o = lds64_if(j < c && j != own, s_keys + b0 + j);
          r += (o < me) ? 1u : 0u;
        }
        for (uint32_t j = 4; j < c; ++j) r += (s_keys[b0 + j] < me) ? 1u : 0u;
        slot[k] = (uint16_t)(b0 + r);
      }
    }
  } else {
    // generic path (a window that spans nearly the whole key range): full key, then position
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const uint32_t i = tid + k * THREADS;
      if (i < len) {
        const uint32_t d = (uint32_t)((key[k] - base) >> s);
        const uint32_t p = s_bin[d] + slot[k];
        s_keys[p] = key[k];
        s_idx[p] = (uint16_t)i;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const uint32_t i = tid + k * THREADS;
      if (i < len) {
        const uint32_t d = (uint32_t)((key[k] - base) >> s);
        const uint32_t b0 = s_bin[d], b1 = s_bin[d + 1];
        uint32_t r = 0;
        for (uint32_t j = b0; j < b1; ++j) {
          const uint64_t o = s_keys[j];
          r += (o < key[k] || (o == key[k] && s_idx[j] < (uint16_t)i)) ? 1u : 0u;
        }
        slot[k] = (uint16_t)(b0 + r);
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const uint32_t i = tid + k * THREADS;
    if (i < len) {
      s_keys[slot[k]] = key[k];
      s_vals[slot[k]] = val[k];
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const uint32_t i = tid + k * THREADS;
    if (i < len) {
      const uint64_t kk = s_keys[i];
      const uint32_t v = s_vals[i];
      uint32_t at = i;
      if (PADDED) {
        // equal keys arrived in no particular order: a run of them is ordered by row id
        const bool left = i > 0 && s_keys[i - 1] == kk, right = i + 1 < len && s_keys[i + 1] == kk;
        if (left || right) {
          uint32_t a = i, b = i + 1, r = 0;
          while (a > 0 && s_keys[a - 1] == kk) {
            --a;
            r += s_vals[a] < v ? 1u : 0u;
          }
          while (b < len && s_keys[b] == kk) {
            r += s_vals[b] < v ? 1u : 0u;
            ++b;
          }
          at = a + r;
        }
      }
      out_keys[out_start + i] = kk;
      out_vals[out_start + at] = v;
    }
  }
}

// win[0..nwin]: window w is [win[w], win[w+1]).
//   mode A (stride == 0): window w = the segments of prefix values [w << gshift, (w+1) << gshift),
//   nwin = 2^(q - gshift).
//   mode B (stride  > 0): window w starts at the first segment boundary at or after w*stride;
//   the search is bounded by `reach` pairs (a longer segment overflows the window anyway).
__global__ void segment_windows_kernel(const uint64_t* __restrict__ keys, uint32_t n, int q,
                                       uint32_t stride, int gshift, uint32_t reach,
                                       uint32_t nwin, uint32_t* __restrict__ win) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w > nwin) return;
  if (w == nwin) {
    win[w] = n;
    return;
  }
  if (w == 0 || q == 0) {
    win[w] = w == 0 ? 0 : n;
    return;
  }
  const int pshift = 64 - q;
  uint32_t lo, hi;    // answer in [lo, hi]
  uint64_t target;    // first position whose prefix is >= target
  if (stride == 0) {
    lo = 0;
    hi = n;
    target = (uint64_t)w << gshift;
  } else {
    const uint64_t pos64 = (uint64_t)w * stride;
    if (pos64 >= n) {
      win[w] = n;
      return;
    }
    const uint32_t pos = (uint32_t)pos64;
    const uint64_t p = keys[pos - 1] >> pshift;
    if ((keys[pos] >> pshift) != p) {
      win[w] = pos;
      return;
    }
    target = p + 1;
    lo = pos + 1;
    hi = (uint32_t)min((uint64_t)n, (uint64_t)pos + reach);
  }
  while (lo < hi) {
    const uint32_t mid = lo + (hi - lo) / 2;
    if ((keys[mid] >> pshift) >= target) hi = mid;
    else lo = mid + 1;
  }
  win[w] = lo;
}

using SegDefault = SegCfg<512, 9>;

}  // namespace

int hmj_segment_sort_cap() { return SegDefault::CAP; }

// Sorts (keys, vals), already ordered on key >> (64 - q), in place.  group_shift >= 0 selects
// mode A (windows of 2^group_shift prefix values; needs q - group_shift <= 20 and every window
// <= CAP), group_shift < 0 mode B with windows of `stride` pairs.  *d_flag is OR-ed with 1 when
// a window or bin was too large.
int hmj_launch_segment_sort(hmj_ctx* ctx, uint64_t* keys, uint32_t* vals, uint32_t n, int q,
                            int group_shift, uint32_t stride, uint32_t* d_flag) {
  HMJ_ARG(ctx, q >= 0 && q <= 56 && (group_shift < 0 || (group_shift <= q && q - group_shift <= 20)));
  cudaStream_t st = ctx->stream;
  uint32_t nwin;
  if (q == 0) {
    nwin = 1;
    stride = n;
    group_shift = 0;
  } else if (group_shift >= 0) {
    nwin = 1u << (q - group_shift);
    stride = 0;
  } else {
    HMJ_ARG(ctx, stride > 0);
    nwin = (uint32_t)hmj_div_up(n, stride);
    group_shift = 0;
  }
  uint32_t* d_win = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_SEG_WINDOWS, (size_t)nwin + 1, &d_win));
  segment_windows_kernel<<<(nwin + 1 + 255) / 256, 256, 0, st>>>(
      keys, n, q, stride, group_shift, (uint32_t)SegDefault::CAP + 2, nwin, d_win);
  HMJ_LAUNCHED(ctx);
  auto kern = segment_sort_kernel<512, 9, 2, false>;
  HMJ_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)SegDefault::SMEM_BYTES));
  kern<<<nwin, 512, SegDefault::SMEM_BYTES, st>>>(keys, vals, d_win, q, d_flag, nullptr, 0, nullptr,
                                                  nullptr);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

// Finish of the uniform sort: bucket w (counts[w] pairs at slot w * in_cap of the padded
// columns, common q-bit prefix w) is sorted and written to out_* at offsets[w].
int hmj_launch_bucket_finish(hmj_ctx* ctx, uint64_t* keys, uint32_t* vals, const uint32_t* counts,
                             const uint32_t* offsets, uint32_t buckets, uint32_t in_cap, int q,
                             uint64_t* out_keys, uint32_t* out_vals, uint32_t* d_flag) {
  HMJ_ARG(ctx, q >= 0 && q <= 32 && in_cap <= (uint32_t)SegDefault::CAP);
  auto kern = segment_sort_kernel<512, 9, 2, true>;
  HMJ_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)SegDefault::SMEM_BYTES));
  kern<<<buckets, 512, SegDefault::SMEM_BYTES, ctx->stream>>>(keys, vals, offsets, q, d_flag, counts,
                                                              in_cap, out_keys, out_vals);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}
