// strings.cu -- moving string keys between rows: the gather that puts a rank's keys into send
// order for the hash-range exchange of the multi-GPU join ("string keys travel with
// their pairs").  The reference moves std::string objects inside its tuples on every scatter
// (/root/reference radix_hash.h:340-345, partitioned_hash.h:102-110); here keys live in an
// offsets[n+1] + chars[] column pair and only this kernel ever copies their bytes.
#include "common.cuh"

namespace {

constexpr int GS_THREADS = 256;
constexpr int GS_LANES = 8;  // threads that share one row

// out_chars[out_offsets[i] ...] = chars[offsets[rows[i]] ... offsets[rows[i]+1])
__global__ void __launch_bounds__(GS_THREADS)
gather_strings_kernel(const uint64_t* __restrict__ offsets, const uint8_t* __restrict__ chars,
                      const uint32_t* __restrict__ rows, size_t n,
                      const uint64_t* __restrict__ out_offsets, uint8_t* __restrict__ out_chars) {
  const size_t gid = (size_t)blockIdx.x * GS_THREADS + threadIdx.x;
  const size_t i = gid / GS_LANES;
  const int lane = (int)(gid % GS_LANES);
  if (i >= n) return;
  const uint32_t row = rows[i];
  const uint64_t b = offsets[row], len = offsets[row + 1] - b, o = out_offsets[i];
  const uint8_t* src = chars + b;
  uint8_t* dst = out_chars + o;
  // 8-byte words where both sides allow it, bytes otherwise (keys average 18 bytes)
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 7) == 0) {
    const uint64_t words = len >> 3;
    for (uint64_t w = lane; w < words; w += GS_LANES)
      reinterpret_cast<uint64_t*>(dst)[w] = reinterpret_cast<const uint64_t*>(src)[w];
    for (uint64_t k = (words << 3) + lane; k < len; k += GS_LANES) dst[k] = src[k];
  } else {
    for (uint64_t k = lane; k < len; k += GS_LANES) dst[k] = src[k];
  }
}

// start[i] = offsets[rows[i]], len[i] = length of that row: ONE random read per row (both offsets
// of a row share a 16-byte neighbourhood), after which the gather below needs no offsets column
__global__ void __launch_bounds__(256)
row_spans_kernel(const uint64_t* __restrict__ offsets, const uint32_t* __restrict__ rows, size_t n,
                 uint64_t* __restrict__ start, uint32_t* __restrict__ len) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t row = rows[i];
    const uint64_t b = offsets[row], e = offsets[row + 1];
    start[i] = b;
    len[i] = (uint32_t)(e - b);
  }
}

// out_chars[out_offsets[i] ...] = chars[start[i] ... start[i] + (out_offsets[i+1] - out_offsets[i]))
__global__ void __launch_bounds__(GS_THREADS)
gather_spans_kernel(const uint8_t* __restrict__ chars, const uint64_t* __restrict__ start, size_t n,
                    const uint64_t* __restrict__ out_offsets, uint8_t* __restrict__ out_chars) {
  const size_t gid = (size_t)blockIdx.x * GS_THREADS + threadIdx.x;
  const size_t i = gid / GS_LANES;
  const int lane = (int)(gid % GS_LANES);
  if (i >= n) return;
  const uint64_t o = out_offsets[i], len = out_offsets[i + 1] - o;
  const uint8_t* src = chars + start[i];
  uint8_t* dst = out_chars + o;
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 7) == 0) {
    const uint64_t words = len >> 3;
    for (uint64_t w = lane; w < words; w += GS_LANES)
      reinterpret_cast<uint64_t*>(dst)[w] = reinterpret_cast<const uint64_t*>(src)[w];
    for (uint64_t k = (words << 3) + lane; k < len; k += GS_LANES) dst[k] = src[k];
  } else {
    for (uint64_t k = lane; k < len; k += GS_LANES) dst[k] = src[k];
  }
}

}  // namespace

extern "C" int hmj_string_row_spans(hmj_ctx* ctx, const uint64_t* d_offsets, const uint32_t* d_rows,
                                    size_t n, uint64_t* d_start_out, uint32_t* d_len_out) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_offsets && d_rows && d_start_out && d_len_out);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  size_t blocks = hmj_div_up(n, 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  row_spans_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(d_offsets, d_rows, n, d_start_out, d_len_out);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

extern "C" int hmj_gather_spans(hmj_ctx* ctx, const uint8_t* d_chars, const uint64_t* d_start, size_t n,
                                const uint64_t* d_out_offsets, uint8_t* d_out_chars) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_start && d_out_offsets);
  HMJ_ARG(ctx, n < (size_t(1) << 32));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t threads = n * GS_LANES;
  gather_spans_kernel<<<(unsigned)hmj_div_up(threads, GS_THREADS), GS_THREADS, 0, ctx->stream>>>(
      d_chars, d_start, n, d_out_offsets, d_out_chars);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

extern "C" int hmj_gather_strings(hmj_ctx* ctx, const uint64_t* d_offsets, const uint8_t* d_chars,
                                  const uint32_t* d_rows, size_t n, const uint64_t* d_out_offsets,
                                  uint8_t* d_out_chars) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_offsets && d_rows && d_out_offsets);
  HMJ_ARG(ctx, n < (size_t(1) << 32));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t threads = n * GS_LANES;
  gather_strings_kernel<<<(unsigned)hmj_div_up(threads, GS_THREADS), GS_THREADS, 0, ctx->stream>>>(
      d_offsets, d_chars, d_rows, n, d_out_offsets, d_out_chars);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}
