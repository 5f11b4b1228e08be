// host_api.cu -- host-buffer flavours of the C ABI: H2D staging, the device path, D2H of the
// result.  These are what the C++ facade (hashmergejoin_b200/facade/*.h) and the end-to-end
// benchmark call; the copies are inside the call on purpose.
#include "common.cuh"

#include <mutex>

namespace {
enum { EV_START = 0, EV_HASHED, EV_SORTED, EV_JOINED, EV_H2D, EV_D2H };

// One host thread at a time per GPU and direction on the host link.  Large copies issued from
// several streams share a copy engine slice by slice, so joins in flight from several host
// threads (a context each) would all finish their input copies together and then all copy
// their pairs out together; taking turns lets one join's pairs leave while the next join's
// keys arrive (the link is full duplex: 45 + 45 GB/s measured against 55 one way).
std::mutex g_turn[2][64];
class CopyTurn {
 public:
  CopyTurn(hmj_ctx* ctx, int dir) : ctx_(ctx), lock_(g_turn[dir][ctx->device & 63]) {}
  // the copies issued in this turn have landed
  int finish() {
    const cudaError_t e = cudaStreamSynchronize(ctx_->stream);
    lock_.unlock();
    HMJ_CUDA(ctx_, e);
    return HMJ_OK;
  }
 private:
  hmj_ctx* ctx_;
  std::unique_lock<std::mutex> lock_;
};
enum { TURN_H2D = 0, TURN_D2H = 1 };

void mark(hmj_ctx* ctx, const char* what) { hmj_host_mark(ctx, what); }

template <typename T>
int stage_in(hmj_ctx* ctx, int slot, const T* h_src, size_t count, T** d_dst) {
  HMJ_TRY(hmj_ws(ctx, slot, count, d_dst));
  if (count)
    HMJ_CUDA(ctx, cudaMemcpyAsync(*d_dst, h_src, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  return HMJ_OK;
}
template <typename T>
int copy_out(hmj_ctx* ctx, T* h_dst, const T* d_src, size_t count) {
  if (count && h_dst)
    HMJ_CUDA(ctx, cudaMemcpyAsync(h_dst, d_src, count * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  return HMJ_OK;
}
}  // namespace

extern "C" int hmj_join_u64_host(hmj_ctx* ctx, const uint64_t* h_rkeys, size_t nr,
                                 const uint64_t* h_skeys, size_t ns, uint32_t* h_out_r,
                                 uint32_t* h_out_s, size_t cap, uint64_t* h_num_pairs) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, h_num_pairs && (nr == 0 || h_rkeys) && (ns == 0 || h_skeys));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  uint64_t *dr = nullptr, *ds = nullptr;
  mark(ctx, "enter");
  {
    CopyTurn turn(ctx, TURN_H2D);
    mark(ctx, "h2d turn");
    HMJ_TRY(stage_in(ctx, WS_IN_0, h_rkeys, nr, &dr));
    HMJ_TRY(stage_in(ctx, WS_IN_1, h_skeys, ns, &ds));
    HMJ_TRY(turn.finish());
  }
  mark(ctx, "h2d done");
  HMJ_TRY(hmj_join_u64(ctx, dr, nr, ds, ns, h_num_pairs));
  mark(ctx, "joined");
  if (h_out_r || h_out_s) {
    CopyTurn turn(ctx, TURN_D2H);
    mark(ctx, "d2h turn");
    HMJ_TRY(hmj_join_result_to_host(ctx, h_out_r, h_out_s, cap));
  }
  mark(ctx, "leave");
  return HMJ_OK;
}

extern "C" int hmj_join_strings_host(hmj_ctx* ctx, const uint64_t* h_roffsets,
                                     const uint8_t* h_rchars, size_t nr,
                                     const uint64_t* h_soffsets, const uint8_t* h_schars,
                                     size_t ns, uint32_t* h_out_r, uint32_t* h_out_s, size_t cap,
                                     uint64_t* h_num_pairs) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, h_num_pairs && h_roffsets && h_soffsets);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  uint64_t *dro = nullptr, *dso = nullptr;
  uint8_t *drc = nullptr, *dsc = nullptr;
  {
    CopyTurn turn(ctx, TURN_H2D);
    HMJ_TRY(stage_in(ctx, WS_IN_0, h_roffsets, nr + 1, &dro));
    HMJ_TRY(stage_in(ctx, WS_IN_1, h_rchars, (size_t)h_roffsets[nr], &drc));
    HMJ_TRY(stage_in(ctx, WS_IN_2, h_soffsets, ns + 1, &dso));
    HMJ_TRY(stage_in(ctx, WS_IN_3, h_schars, (size_t)h_soffsets[ns], &dsc));
    HMJ_TRY(turn.finish());
  }
  HMJ_TRY(hmj_join_strings(ctx, dro, drc, nr, dso, dsc, ns, h_num_pairs));
  if (h_out_r || h_out_s) {
    CopyTurn turn(ctx, TURN_D2H);
    HMJ_TRY(hmj_join_result_to_host(ctx, h_out_r, h_out_s, cap));
  }
  return HMJ_OK;
}

extern "C" int hmj_hash_sort_u64_host(hmj_ctx* ctx, const uint64_t* h_keys, size_t n,
                                      uint64_t* h_hash_out, uint32_t* h_rowid_out) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, h_keys != nullptr);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  uint64_t *dk = nullptr, *dh = nullptr;
  uint32_t* dr = nullptr;
  HMJ_TRY(stage_in(ctx, WS_IN_0, h_keys, n, &dk));
  HMJ_TRY(hmj_ws(ctx, WS_R_HASH, n, &dh));
  HMJ_TRY(hmj_ws(ctx, WS_R_ROWID, n, &dr));
  HMJ_TRY(hmj_hash_sort_u64(ctx, dk, n, dh, dr));
  HMJ_TRY(copy_out(ctx, h_hash_out, dh, n));
  HMJ_TRY(copy_out(ctx, h_rowid_out, dr, n));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_hash_sort_strings_host(hmj_ctx* ctx, const uint64_t* h_offsets,
                                          const uint8_t* h_chars, size_t n, uint64_t* h_hash_out,
                                          uint32_t* h_rowid_out) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, h_offsets != nullptr);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  uint64_t *doff = nullptr, *dh = nullptr;
  uint8_t* dc = nullptr;
  uint32_t* dr = nullptr;
  HMJ_TRY(stage_in(ctx, WS_IN_0, h_offsets, n + 1, &doff));
  HMJ_TRY(stage_in(ctx, WS_IN_1, h_chars, (size_t)h_offsets[n], &dc));
  HMJ_TRY(hmj_ws(ctx, WS_R_HASH, n, &dh));
  HMJ_TRY(hmj_ws(ctx, WS_R_ROWID, n, &dr));
  HMJ_TRY(hmj_hash_sort_strings(ctx, doff, dc, n, dh, dr));
  HMJ_TRY(copy_out(ctx, h_hash_out, dh, n));
  HMJ_TRY(copy_out(ctx, h_rowid_out, dr, n));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_sort_pairs_host(hmj_ctx* ctx, const uint64_t* h_hash, const uint32_t* h_rowid,
                                   size_t n, uint64_t* h_hash_out, uint32_t* h_rowid_out) {
  if (!ctx) return HMJ_EARG;
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, h_hash != nullptr);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  uint64_t *dk = nullptr, *dh = nullptr;
  uint32_t *dv = nullptr, *dr = nullptr;
  HMJ_TRY(stage_in(ctx, WS_IN_0, h_hash, n, &dk));
  if (h_rowid) HMJ_TRY(stage_in(ctx, WS_IN_1, h_rowid, n, &dv));
  HMJ_TRY(hmj_ws(ctx, WS_R_HASH, n, &dh));
  HMJ_TRY(hmj_ws(ctx, WS_R_ROWID, n, &dr));
  HMJ_TRY(hmj_sort_pairs(ctx, dk, dv, n, dh, dr));
  HMJ_TRY(copy_out(ctx, h_hash_out, dh, n));
  HMJ_TRY(copy_out(ctx, h_rowid_out, dr, n));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_partition_counts_host(hmj_ctx* ctx, const uint64_t* h_hash, size_t n, int bits,
                                         uint64_t* h_counts) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, bits >= 1 && bits <= 16 && h_counts && (n == 0 || h_hash));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  uint64_t *dk = nullptr, *dc = nullptr;
  HMJ_TRY(stage_in(ctx, WS_IN_0, h_hash, n, &dk));
  HMJ_TRY(hmj_ws(ctx, WS_IN_1, size_t(1) << bits, &dc));
  HMJ_TRY(hmj_partition_counts(ctx, dk, n, bits, dc));
  HMJ_TRY(copy_out(ctx, h_counts, dc, size_t(1) << bits));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HMJ_OK;
}
