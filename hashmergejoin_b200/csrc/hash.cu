// hash.cu -- K1: 64-bit key hashing, bit-exact with libstdc++ std::hash on x86-64.
//
// Stands in for `Hash{}(std::get<0>(*iter))` at /root/reference radix_hash.h:314,340 and
// partitioned_hash.h:58,76 with the default Hash = std::hash<Key>:
//   * integral keys: identity (libstdc++ functional_hash.h).
//   * std::string : std::_Hash_bytes(data, len, 0xc70f6907) -- a Murmur64A-style hash whose
//     source is not in the reference tree (libstdc++ 13.3, hash_bytes.cc); the published
//     algorithm is restated in hash_device.cuh and pinned by tests/golden/std_hash_strings.json.
//
// Layout: keys are columnar -- offsets u64[n+1] + chars u8[].  A CTA takes a tile of
// consecutive rows, whose bytes are one contiguous span of `chars`; the span is staged into
// shared memory with coalesced 16-byte loads and each thread then hashes whole rows out of
// shared memory with aligned 8-byte reads + funnel shifts (rows start at arbitrary byte
// offsets).  Algorithmic bytes per row: L + 8 (offset) read, 8 written.
#include "common.cuh"
#include "hash_device.cuh"

namespace {

constexpr int HASH_THREADS = 256;
constexpr int HASH_ROWS_PER_THREAD = 4;
constexpr int HASH_TILE_ROWS = HASH_THREADS * HASH_ROWS_PER_THREAD;  // 1024 rows
constexpr int HASH_SMEM_CHARS = 36 * 1024;                           // staged bytes per tile

__global__ void __launch_bounds__(HASH_THREADS)
hash_strings_kernel(const uint64_t* __restrict__ offsets, const uint8_t* __restrict__ chars,
                    size_t n, uint64_t* __restrict__ out) {
  __shared__ uint64_t s_off[HASH_TILE_ROWS + 1];
  __shared__ __align__(16) uint8_t s_chars[HASH_SMEM_CHARS + 32];

  const size_t num_tiles = (n + HASH_TILE_ROWS - 1) / HASH_TILE_ROWS;
  for (size_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const size_t r0 = tile * HASH_TILE_ROWS;
    const int rows = (int)min((size_t)HASH_TILE_ROWS, n - r0);
    __syncthreads();  // previous tile's readers are done with smem
    for (int i = threadIdx.x; i <= rows; i += HASH_THREADS) s_off[i] = offsets[r0 + i];
    __syncthreads();
    const uint64_t c0 = s_off[0], c1 = s_off[rows];
    // absolute address of the first byte, aligned down to 16 for vector loads
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(chars) + c0;
    const uintptr_t a0_al = a0 & ~uintptr_t(15);
    const uint32_t lead = (uint32_t)(a0 - a0_al);
    const uint64_t span = (c1 - c0) + lead;
    const bool staged = span <= (uint64_t)HASH_SMEM_CHARS;
    if (staged) {
      const uint32_t nvec = (uint32_t)((span + 15) >> 4);
      const uint4* src = reinterpret_cast<const uint4*>(a0_al);
      uint4* dst = reinterpret_cast<uint4*>(s_chars);
      for (uint32_t i = threadIdx.x; i < nvec; i += HASH_THREADS) dst[i] = __ldg(src + i);
      __syncthreads();
      const uint64_t* s_words = reinterpret_cast<const uint64_t*>(s_chars);
      for (int i = threadIdx.x; i < rows; i += HASH_THREADS) {
        const uint64_t b = s_off[i] - c0 + lead;
        const uint32_t len = (uint32_t)(s_off[i + 1] - s_off[i]);
        out[r0 + i] = hmj_dev::hash_bytes_words<false>(s_words, b, len, 0);
      }
    } else {
      // long rows: hash straight from global memory (aligned words, guarded at the end)
      const uintptr_t base_al = reinterpret_cast<uintptr_t>(chars) & ~uintptr_t(7);
      const uint32_t base_lead = (uint32_t)(reinterpret_cast<uintptr_t>(chars) - base_al);
      const uint64_t* g_words = reinterpret_cast<const uint64_t*>(base_al);
      for (int i = threadIdx.x; i < rows; i += HASH_THREADS) {
        const uint64_t b = s_off[i] + base_lead;
        const uint64_t len = s_off[i + 1] - s_off[i];
        out[r0 + i] = hmj_dev::hash_bytes_words<true>(g_words, b, len, 0);
      }
    }
  }
}

// ---- bulk-copy pipelined variant ---------------------------------------------------------
// The kernel above loads a tile, waits, hashes, and only then turns to the next tile: ncu showed
// 29 % of its stall samples on the STS that stages freshly loaded bytes (profiles/r2j_hash_strings_
// digest.txt).  Here persistent CTAs keep the NEXT tile's offsets and key bytes in flight as two
// bulk async copies (cp.async.bulk, completion on an mbarrier) into a second pair of buffers while
// the current tile is hashed out of shared memory.  A tile whose bytes do not fit the buffer, the
// last tile (its copies would read past the columns) and misaligned columns take the plain-load
// route inside the same kernel.
constexpr int HP_THREADS = 512;
constexpr int HP_TILE_ROWS = 1024;
constexpr int HP_CHARS = 36 * 1024;
constexpr int HP_OFF_BYTES = (HP_TILE_ROWS + 2) * 8;  // 1026 offsets: a multiple of 16 bytes
constexpr size_t HP_SMEM = 2 * (size_t)(HP_CHARS + 32) + 2 * (size_t)HP_OFF_BYTES + 64;

__device__ __forceinline__ uint32_t hp_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(HP_THREADS, 2)
hash_strings_tma_kernel(const uint64_t* __restrict__ offsets, const uint8_t* __restrict__ chars,
                        size_t n, uint64_t* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char hp_raw[];
  unsigned char* s_chars[2] = {hp_raw, hp_raw + HP_CHARS + 32};
  uint64_t* s_off[2] = {reinterpret_cast<uint64_t*>(hp_raw + 2 * (HP_CHARS + 32)),
                        reinterpret_cast<uint64_t*>(hp_raw + 2 * (HP_CHARS + 32) + HP_OFF_BYTES)};
  uint64_t* s_bar = reinterpret_cast<uint64_t*>(hp_raw + 2 * (HP_CHARS + 32) + 2 * HP_OFF_BYTES);  // [2]
  uint32_t* s_issued = reinterpret_cast<uint32_t*>(s_bar + 2);                                    // [2]
  const int tid = threadIdx.x;
  const size_t num_tiles = (n + HP_TILE_ROWS - 1) / HP_TILE_ROWS;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(hp_smem(&s_bar[0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(hp_smem(&s_bar[1])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    s_issued[0] = s_issued[1] = 0;
  }
  __syncthreads();
  // thread 0: start the copies of `tile` into buffer b when the tile is regular
  auto issue = [&](size_t tile, int b) {
    if (tile >= num_tiles) return;
    const size_t r0 = tile * HP_TILE_ROWS;
    if (r0 + HP_TILE_ROWS + 2 > n + 1) return;  // last tile(s): offsets copy would leave the column
    const uint64_t c0 = offsets[r0], c1 = offsets[r0 + HP_TILE_ROWS];
    const uint64_t a0 = c0 & ~uint64_t(15);
    const uint64_t bytes = ((c1 - a0) + 15) & ~uint64_t(15);
    if (bytes > (uint64_t)HP_CHARS || bytes == 0) return;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(hp_smem(&s_bar[b])),
                 "r"((uint32_t)bytes + (uint32_t)HP_OFF_BYTES)
                 : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                     "r"(hp_smem(s_off[b])),
                 "l"(offsets + r0), "r"((uint32_t)HP_OFF_BYTES), "r"(hp_smem(&s_bar[b]))
                 : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                     "r"(hp_smem(s_chars[b])),
                 "l"(chars + a0), "r"((uint32_t)bytes), "r"(hp_smem(&s_bar[b]))
                 : "memory");
    s_issued[b] = 1;
  };
  uint32_t phase[2] = {0, 0};
  if (tid == 0) issue(blockIdx.x, 0);
  __syncthreads();
  int it = 0;
  for (size_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const int b = it & 1;
    const size_t r0 = tile * HP_TILE_ROWS;
    const int rows = (int)min((size_t)HP_TILE_ROWS, n - r0);
    if (tid == 0) issue(tile + gridDim.x, b ^ 1);  // buffer b^1 was released by the previous iteration
    const bool bulk = s_issued[b] != 0;
    if (bulk) {
      asm volatile(
          "{\n\t"
          ".reg .pred p;\n\t"
          "WAIT_%=:\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
          "@p bra DONE_%=;\n\t"
          "bra WAIT_%=;\n\t"
          "DONE_%=:\n\t"
          "}" ::"r"(hp_smem(&s_bar[b])),
          "r"(phase[b])
          : "memory");
      phase[b] ^= 1;
      const uint64_t c0 = s_off[b][0];
      const uint32_t lead = (uint32_t)(c0 & 15);
      const uint64_t* s_words = reinterpret_cast<const uint64_t*>(s_chars[b]);
      for (int i = tid; i < rows; i += HP_THREADS) {
        const uint64_t o = s_off[b][i];
        const uint32_t len = (uint32_t)(s_off[b][i + 1] - o);
        out[r0 + i] = hmj_dev::hash_bytes_words<false>(s_words, o - c0 + lead, len, 0);
      }
    } else {
      // irregular tile: offsets by plain loads, key bytes hashed straight from global memory
      const uintptr_t base_al = reinterpret_cast<uintptr_t>(chars) & ~uintptr_t(7);
      const uint32_t base_lead = (uint32_t)(reinterpret_cast<uintptr_t>(chars) - base_al);
      const uint64_t* g_words = reinterpret_cast<const uint64_t*>(base_al);
      for (int i = tid; i < rows; i += HP_THREADS) {
        const uint64_t o = offsets[r0 + i], len = offsets[r0 + i + 1] - o;
        out[r0 + i] = hmj_dev::hash_bytes_words<true>(g_words, o + base_lead, len, 0);
      }
    }
    __syncthreads();  // everybody is done with buffer b and has read s_issued[b]
    if (tid == 0) s_issued[b] = 0;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
hash_u64_kernel(const uint64_t* __restrict__ keys, size_t n, uint64_t* __restrict__ out) {
  // std::hash<integral> is the identity; vectorised 16-byte copy.
  const size_t n2 = n >> 1;
  const ulonglong2* src = reinterpret_cast<const ulonglong2*>(keys);
  ulonglong2* dst = reinterpret_cast<ulonglong2*>(out);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride)
    dst[i] = src[i];
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) out[n - 1] = keys[n - 1];
}

__global__ void hash_u64_kernel_unaligned(const uint64_t* __restrict__ keys, size_t n,
                                          uint64_t* __restrict__ out) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = keys[i];
}

}  // namespace

extern "C" int hmj_hash_strings(hmj_ctx* ctx, const uint64_t* d_offsets, const uint8_t* d_chars,
                                size_t n, uint64_t* d_hash) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_offsets && d_hash);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t tiles = hmj_div_up(n, HASH_TILE_ROWS);
  const int grid = (int)min(tiles, (size_t)ctx->num_sms * 16);
  cudaEvent_t ev;  // 8 (offset) + 8 (hash) per row; the key bytes are added by whoever knows them
  HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_HASH, (double)n * 16.0, &ev));
  // bulk copies need 16-byte aligned columns (HMJ_HASH_TMA=0: the plain-load kernel)
  static const bool use_tma = [] {
    const char* e = getenv("HMJ_HASH_TMA");
    return !(e && atoi(e) == 0);  // C2: 0.666 -> 0.623 ms per 64M rows
  }();
  if (use_tma && ((reinterpret_cast<uintptr_t>(d_offsets) | reinterpret_cast<uintptr_t>(d_chars)) & 15) == 0 &&
      tiles >= 4) {
    HMJ_CUDA(ctx, cudaFuncSetAttribute(hash_strings_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)HP_SMEM));
    const int g2 = (int)min(tiles, (size_t)ctx->num_sms * 2);
    hash_strings_tma_kernel<<<g2, HP_THREADS, HP_SMEM, ctx->stream>>>(d_offsets, d_chars, n, d_hash);
  } else {
    hash_strings_kernel<<<grid, HASH_THREADS, 0, ctx->stream>>>(d_offsets, d_chars, n, d_hash);
  }
  HMJ_LAUNCHED(ctx);
  hmj_prof_end(ctx, ev);
  return HMJ_OK;
}

extern "C" int hmj_hash_u64(hmj_ctx* ctx, const uint64_t* d_keys, size_t n, uint64_t* d_hash) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_keys && d_hash);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  const int grid = (int)min(hmj_div_up(n, 256 * 8), (size_t)ctx->num_sms * 8);
  const bool aligned = ((reinterpret_cast<uintptr_t>(d_keys) | reinterpret_cast<uintptr_t>(d_hash)) & 15) == 0;
  if (aligned)
    hash_u64_kernel<<<grid, 256, 0, ctx->stream>>>(d_keys, n, d_hash);
  else
    hash_u64_kernel_unaligned<<<grid, 256, 0, ctx->stream>>>(d_keys, n, d_hash);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}
