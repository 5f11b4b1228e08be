// context.cu -- context lifetime, workspace and small host helpers of the C ABI (include/hmj.h).
#include <math.h>

#include <new>

#include "common.cuh"

#include <chrono>
#include <cstdlib>

extern "C" const char* hmj_version(void) { return "hmj-b200 0.1 sm_100a"; }

static int hmj_create_impl(hmj_ctx** out, int device, void* stream, bool own) {
  if (!out) return HMJ_EARG;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) return HMJ_ENODEVICE;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return HMJ_ENODEVICE;
  // the kernels are compiled for sm_100a only; anything else cannot run them.
  if (prop.major != 10) return HMJ_ENODEVICE;
  hmj_ctx* ctx = new (std::nothrow) hmj_ctx();
  if (!ctx) return HMJ_ENOMEM;
  ctx->device = device;
  ctx->num_sms = prop.multiProcessorCount;
  memset(&ctx->tm, 0, sizeof ctx->tm);
  if (cudaSetDevice(device) != cudaSuccess) {
    delete ctx;
    return HMJ_ECUDA;
  }
  if (own) {
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
      delete ctx;
      return HMJ_ECUDA;
    }
    ctx->own_stream = true;
  } else {
    ctx->stream = reinterpret_cast<cudaStream_t>(stream);
  }
  ctx->pinned_bytes = 1 << 16;
  if (cudaMallocHost(&ctx->pinned, ctx->pinned_bytes) != cudaSuccess) {
    if (own) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return HMJ_ENOMEM;
  }
  for (auto& ev : ctx->ev) cudaEventCreate(&ev);
  *out = ctx;
  return HMJ_OK;
}

extern "C" int hmj_create(hmj_ctx** ctx, int device) {
  return hmj_create_impl(ctx, device, nullptr, true);
}

extern "C" int hmj_create_on_stream(hmj_ctx** ctx, int device, void* cuda_stream) {
  return hmj_create_impl(ctx, device, cuda_stream, false);
}

extern "C" int hmj_release_workspace(hmj_ctx* ctx) {
  if (!ctx) return HMJ_EARG;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (auto& b : ctx->ws) {
    if (b.ptr) cudaFree(b.ptr);
    b.ptr = nullptr;
    b.bytes = 0;
  }
  // the look-back descriptors are gone with WS_LOOKBACK: the next sort must clear its buffer even
  // if cudaMalloc hands back the same address (stale words could carry a live epoch tag)
  ctx->lb_seen_ptr = nullptr;
  ctx->lb_seen_bytes = 0;
  ctx->lb_epoch = 0;
  return HMJ_OK;
}

extern "C" int hmj_dist_teardown(hmj_ctx* ctx);
extern "C" int hmj_xchg_teardown(hmj_ctx* ctx);

extern "C" int hmj_destroy(hmj_ctx* ctx) {
  if (!ctx) return HMJ_EARG;
  hmj_dist_teardown(ctx);
  hmj_xchg_teardown(ctx);
  hmj_release_workspace(ctx);
  for (auto& ev : ctx->ev)
    if (ev) cudaEventDestroy(ev);
  for (auto& ev : ctx->prof_events) cudaEventDestroy(ev);
  for (auto& ev : ctx->copy_events) cudaEventDestroy(ev);
  if (!ctx->sample_host.empty()) cudaHostUnregister(ctx->sample_host.data());
  if (ctx->copy_in) cudaStreamDestroy(ctx->copy_in);
  if (ctx->copy_out) cudaStreamDestroy(ctx->copy_out);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return HMJ_OK;
}

extern "C" const char* hmj_last_error(hmj_ctx* ctx) {
  if (!ctx) return "null context";
  return ctx->err.c_str();
}

extern "C" int hmj_synchronize(hmj_ctx* ctx) {
  if (!ctx) return HMJ_EARG;
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_launch_count(hmj_ctx* ctx, uint64_t* count) {
  if (!ctx || !count) return HMJ_EARG;
  *count = ctx->launches;
  return HMJ_OK;
}

namespace {
__global__ void __launch_bounds__(256) fill_words_kernel(uint32_t* __restrict__ p, uint32_t word, size_t words) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
    uint4* q = reinterpret_cast<uint4*>(p);
    const uint4 v = make_uint4(word, word, word, word);
    for (size_t k = i; k < words / 4; k += stride) q[k] = v;
    for (size_t k = (words & ~size_t(3)) + i; k < words; k += stride) p[k] = word;
  } else {
    for (; i < words; i += stride) p[i] = word;
  }
}
}  // namespace

void hmj_host_mark(hmj_ctx* ctx, const char* what) {
  static const bool on = getenv("HMJ_HOST_TRACE") != nullptr;
  if (!on) return;
  const double ms = std::chrono::duration<double, std::milli>(
                        std::chrono::steady_clock::now().time_since_epoch()).count();
  fprintf(stderr, "[hmj host %p] %.1f %s\n", (void*)ctx, ms, what);
}

int hmj_fill_async(hmj_ctx* ctx, void* ptr, int byte, size_t bytes, cudaStream_t st) {
  if (bytes == 0) return HMJ_OK;
  HMJ_ARG(ctx, ptr != nullptr && (bytes & 3) == 0 && (reinterpret_cast<uintptr_t>(ptr) & 3) == 0);
  const uint32_t word = 0x01010101u * (uint32_t)(byte & 0xFF);
  const size_t words = bytes / 4;
  size_t blocks = (words / 4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  fill_words_kernel<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<uint32_t*>(ptr), word, words);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

int hmj_ws_get(hmj_ctx* ctx, int slot, size_t bytes, void** out) {
  hmj_buf& b = ctx->ws[slot];
  if (bytes == 0) bytes = 16;
  if (b.bytes < bytes) {
    HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
    if (b.ptr) {
      // the old buffer may still be in use by queued work on our stream
      HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      HMJ_CUDA(ctx, cudaFree(b.ptr));
      b.ptr = nullptr;
      b.bytes = 0;
      if (slot == WS_LOOKBACK) {
        ctx->lb_seen_ptr = nullptr;
        ctx->lb_seen_bytes = 0;
        ctx->lb_epoch = 0;
      }
    }
    size_t want = (bytes + 255) & ~size_t(255);
    HMJ_CUDA(ctx, cudaMalloc(&b.ptr, want));
    b.bytes = want;
  }
  *out = b.ptr;
  return HMJ_OK;
}

// Radix width helper kept for API parity with radix_hash::optimal_partition
// (radix_hash.h:38-57): among k = 6..14 pick the k whose log_{2^k}(n) is nearest an integer
// (the last such k on ties); n below 2^k returns that k at once.  The GPU sort itself always
// uses 8-bit digits; this only serves callers that ask for the reference's choice.
extern "C" int hmj_optimal_partition(size_t input_num) {
  double best = 1.0;
  int pick = 0;
  for (int k = 6; k <= 14; ++k) {
    if (input_num < (size_t(1) << k)) return k;
    double x = log((double)input_num) / log((double)(1 << k));
    double frac_lo = x - floor(x), frac_hi = ceil(x) - x;
    double d = frac_lo < frac_hi ? frac_lo : frac_hi;
    if (d <= best) {
      best = d;
      pick = k;
    }
  }
  return pick;
}
