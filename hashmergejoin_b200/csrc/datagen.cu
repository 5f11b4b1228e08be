// datagen.cu -- device-side generators of the benchmark relations.
//
// The reference builds its benchmark inputs on the host: create_strvec (/root/reference
// strgen.cc:28-61) reads a word list, emits ceil(sqrt(n)) base words w_i, then "w_i-w_j" row-major
// until n rows exist, and std::random_shuffle's them; hashjoin_bench.cc:112-113 calls it once per
// relation.  At the sizes of BASELINE.json's string configurations (64M .. 512M rows) doing that
// with numpy takes most of a minute per relation, so the same shapes are produced here on the GPU,
// straight into the offsets[n+1] + chars[] column pair the join takes:
//   hmj_gen_permute          a seeded bijection of [0, domain) (Feistel network + cycle walking):
//                            the shuffle, applied to row numbers instead of to rows
//   hmj_gen_zipf             rank k with P(k) ~ 1/(k+1)^s by inverse CDF: binary search of a
//                            counter-based uniform in a device prefix table (the Zipf probe side
//                            of SURVEY.md 8d C5)
//   hmj_gen_strgen_lengths   byte length of strgen key `id` (id < q: w_id, else w_i-w_j with
//                            (i, j) = divmod(id - q, q))
//   hmj_gen_strgen_write     the keys' bytes at given offsets
// Every function has a numpy restatement in hashmergejoin_b200/datagen.py that the GPU tests
// compare bit for bit at small n.  Harness code: nothing on the join path calls it.
#include "common.cuh"

namespace {

__host__ __device__ __forceinline__ uint32_t fmix32(uint32_t h) {
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {  // splitmix64 finaliser
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// 4-round balanced Feistel network on 2 * half_bits bits, walked until the value falls inside
// [0, domain): a bijection of [0, domain) for every seed.
__device__ __forceinline__ uint32_t feistel(uint32_t x, uint32_t domain, int half_bits, uint32_t seed) {
  const uint32_t hm = (1u << half_bits) - 1u;
  do {
    uint32_t l = x >> half_bits, r = x & hm;
#pragma unroll
    for (uint32_t round = 0; round < 4; ++round) {
      const uint32_t f = fmix32(r ^ (seed + round * 0x9E3779B9u)) & hm;
      const uint32_t nl = r;
      r = l ^ f;
      l = nl;
    }
    x = (l << half_bits) | r;
  } while (x >= domain);
  return x;
}

__global__ void __launch_bounds__(256)
permute_kernel(const uint32_t* __restrict__ in, uint32_t in_base, size_t n, uint32_t domain,
               int half_bits, uint32_t seed, uint32_t* __restrict__ out) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t x = in ? in[i] : in_base + (uint32_t)i;
    out[i] = feistel(x, domain, half_bits, seed);
  }
}

// u in [0, 1) from (seed, row), then the first k with cdf[k] > u (cdf[n_keys - 1] == 1.0)
__global__ void __launch_bounds__(256)
zipf_kernel(const double* __restrict__ cdf, uint32_t n_keys, uint64_t seed, uint64_t row_base,
            size_t n, uint32_t* __restrict__ out) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t bits = mix64(seed * 0xD1342543DE82EF95ull + row_base + i);
    const double u = (double)(bits >> 11) * (1.0 / 9007199254740992.0);
    uint32_t lo = 0, hi = n_keys - 1;  // answer in [lo, hi]
    while (lo < hi) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      if (__ldg(cdf + mid) > u) hi = mid; else lo = mid + 1;
    }
    out[i] = lo;
  }
}

__device__ __forceinline__ void strgen_parts(uint32_t id, uint32_t q, uint32_t* i, uint32_t* j) {
  if (id < q) {
    *i = id;
    *j = 0xFFFFFFFFu;
  } else {
    const uint32_t p = id - q;
    *i = p / q;
    *j = p - *i * q;
  }
}

__global__ void __launch_bounds__(256)
strgen_lengths_kernel(const uint64_t* __restrict__ woff, uint32_t q, const uint32_t* __restrict__ ids,
                      size_t n, uint64_t* __restrict__ out_len) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += stride) {
    uint32_t i, j;
    strgen_parts(ids[r], q, &i, &j);
    uint64_t len = __ldg(woff + i + 1) - __ldg(woff + i);
    if (j != 0xFFFFFFFFu) len += 1 + __ldg(woff + j + 1) - __ldg(woff + j);
    out_len[r] = len;
  }
}

// 8 threads per row; the word table (a few hundred KB) stays in L1/L2
__global__ void __launch_bounds__(256)
strgen_write_kernel(const uint64_t* __restrict__ woff, const uint8_t* __restrict__ wchars, uint32_t q,
                    const uint32_t* __restrict__ ids, size_t n, const uint64_t* __restrict__ offsets,
                    uint8_t* __restrict__ chars) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t r = gid >> 3;
  const uint32_t lane = (uint32_t)(gid & 7);
  if (r >= n) return;
  uint32_t i, j;
  strgen_parts(ids[r], q, &i, &j);
  const uint64_t wi = __ldg(woff + i);
  const uint32_t li = (uint32_t)(__ldg(woff + i + 1) - wi);
  uint8_t* dst = chars + offsets[r];
  for (uint32_t k = lane; k < li; k += 8) dst[k] = __ldg(wchars + wi + k);
  if (j != 0xFFFFFFFFu) {
    const uint64_t wj = __ldg(woff + j);
    const uint32_t lj = (uint32_t)(__ldg(woff + j + 1) - wj);
    if (lane == 0) dst[li] = (uint8_t)'-';
    for (uint32_t k = lane; k < lj; k += 8) dst[li + 1 + k] = __ldg(wchars + wj + k);
  }
}

unsigned grid_for(size_t n, int per_block) {
  size_t b = (n + per_block - 1) / per_block;
  if (b > 148 * 16) b = 148 * 16;
  return (unsigned)(b ? b : 1);
}

}  // namespace

extern "C" int hmj_gen_permute(hmj_ctx* ctx, const uint32_t* d_in, uint32_t in_base, size_t n,
                               uint32_t domain, uint32_t seed, uint32_t* d_out) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_out != nullptr && domain >= 1);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  int half_bits = 1;
  while (half_bits < 16 && ((uint64_t)1 << (2 * half_bits)) < (uint64_t)domain) ++half_bits;
  permute_kernel<<<grid_for(n, 256), 256, 0, ctx->stream>>>(d_in, in_base, n, domain, half_bits, seed, d_out);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

extern "C" int hmj_gen_zipf(hmj_ctx* ctx, const double* d_cdf, uint32_t n_keys, uint64_t seed,
                            uint64_t row_base, size_t n, uint32_t* d_out_rank) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_cdf != nullptr && d_out_rank != nullptr && n_keys >= 1);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  zipf_kernel<<<grid_for(n, 256), 256, 0, ctx->stream>>>(d_cdf, n_keys, seed, row_base, n, d_out_rank);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

extern "C" int hmj_gen_strgen_lengths(hmj_ctx* ctx, const uint64_t* d_word_offsets, uint32_t q,
                                      const uint32_t* d_key_ids, size_t n, uint64_t* d_out_len) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_word_offsets && d_key_ids && d_out_len && q >= 1);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  strgen_lengths_kernel<<<grid_for(n, 256), 256, 0, ctx->stream>>>(d_word_offsets, q, d_key_ids, n, d_out_len);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

extern "C" int hmj_gen_strgen_write(hmj_ctx* ctx, const uint64_t* d_word_offsets,
                                    const uint8_t* d_word_chars, uint32_t q,
                                    const uint32_t* d_key_ids, size_t n,
                                    const uint64_t* d_offsets, uint8_t* d_chars) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_word_offsets && d_word_chars && d_key_ids && d_offsets && d_chars && q >= 1);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t threads = n * 8;
  strgen_write_kernel<<<(unsigned)hmj_div_up(threads, 256), 256, 0, ctx->stream>>>(
      d_word_offsets, d_word_chars, q, d_key_ids, n, d_offsets, d_chars);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}
