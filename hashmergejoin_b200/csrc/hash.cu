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
  hash_strings_kernel<<<grid, HASH_THREADS, 0, ctx->stream>>>(d_offsets, d_chars, n, d_hash);
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
