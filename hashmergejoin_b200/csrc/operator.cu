// operator.cu -- whole-operator entry points: HashMergeJoin(r_begin,r_end,s_begin,s_end) as one
// call (hash both relations -> sort both -> merge-join), device- and host-buffer flavours.
// Stands in for the HashMergeJoin constructor + full iteration (/root/reference hashjoin.h:56-68,
// 183-195 as driven by hashjoin_bench.cc:126-133) and for radix_non_inplace_par called from
// host containers (radix_hash.h:351-423).
#include "common.cuh"

namespace {

enum { EV_START = 0, EV_HASHED, EV_SORTED, EV_JOINED, EV_H2D, EV_D2H };

int sort_side_u64(hmj_ctx* ctx, const uint64_t* d_keys, size_t n, int slot_h, int slot_r,
                  uint64_t** d_hash, uint32_t** d_rowid) {
  HMJ_TRY(hmj_ws(ctx, slot_h, n, d_hash));
  HMJ_TRY(hmj_ws(ctx, slot_r, n, d_rowid));
  return hmj_hash_sort_u64(ctx, d_keys, n, *d_hash, *d_rowid);
}

int join_sorted(hmj_ctx* ctx, const hmj_relation* R, const hmj_relation* S, uint64_t* h_num_pairs) {
  return hmj_impl_join_sorted(ctx, R, S, h_num_pairs);
}

}  // namespace

extern "C" int hmj_join_u64(hmj_ctx* ctx, const uint64_t* d_rkeys, size_t nr,
                            const uint64_t* d_skeys, size_t ns, uint64_t* h_num_pairs) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, h_num_pairs != nullptr);
  HMJ_ARG(ctx, (nr == 0 || d_rkeys) && (ns == 0 || d_skeys));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->tm.sort_passes = 0;
  ctx->tm.sort_finishes = 0;
  ctx->tm.sort_fallbacks = 0;
  ctx->tm.sort_uniform = 0;
  ctx->tm.bucket_join = 0;
  ctx->tm.scatter_levels = 0;
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_START], ctx->stream));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_HASHED], ctx->stream));  // identity hash: no kernel
  {
    // uniform hashes: partition both sides on the same hash bits, join bucket by bucket
    int ok = 0;
    HMJ_TRY(hmj_join_uniform_try(ctx, HMJ_KEY_U64, d_rkeys, nullptr, nr, d_skeys, nullptr, ns, 0, nullptr, nullptr,
                                 nullptr, nullptr, ctx->ev[EV_SORTED], h_num_pairs, &ok));
    if (ok) {
      ctx->tm.bucket_join = 1;
      HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_JOINED], ctx->stream));
      return HMJ_OK;
    }
  }
  hmj_relation R = {}, S = {};
  uint64_t* h = nullptr;
  uint32_t* rid = nullptr;
  const bool skew_r = ctx->sample_skewed[0], skew_s = ctx->sample_skewed[1];
  ctx->next_sort_skewed = skew_r;
  HMJ_TRY(sort_side_u64(ctx, d_rkeys, nr, WS_R_HASH, WS_R_ROWID, &h, &rid));
  R.d_hash = h; R.d_rowid = rid; R.n = nr; R.key_kind = HMJ_KEY_U64;
  ctx->next_sort_skewed = skew_s;
  HMJ_TRY(sort_side_u64(ctx, d_skeys, ns, WS_S_HASH, WS_S_ROWID, &h, &rid));
  S.d_hash = h; S.d_rowid = rid; S.n = ns; S.key_kind = HMJ_KEY_U64;
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_SORTED], ctx->stream));
  HMJ_TRY(join_sorted(ctx, &R, &S, h_num_pairs));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_JOINED], ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_join_pairs(hmj_ctx* ctx, const uint64_t* d_rhash, const uint32_t* d_rrowid,
                              size_t nr, const uint64_t* d_shash, const uint32_t* d_srowid,
                              size_t ns, int prefix_bits, uint64_t* h_num_pairs) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, h_num_pairs != nullptr && prefix_bits >= 0 && prefix_bits <= 16);
  HMJ_ARG(ctx, (nr == 0 || (d_rhash && d_rrowid)) && (ns == 0 || (d_shash && d_srowid)));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->tm.sort_passes = 0;
  ctx->tm.sort_finishes = 0;
  ctx->tm.sort_fallbacks = 0;
  ctx->tm.sort_uniform = 0;
  ctx->tm.bucket_join = 0;
  ctx->tm.scatter_levels = 0;
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_START], ctx->stream));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_HASHED], ctx->stream));
  {
    int ok = 0;
    HMJ_TRY(hmj_join_uniform_try(ctx, HMJ_KEY_U64, d_rhash, d_rrowid, nr, d_shash, d_srowid, ns,
                                 prefix_bits, nullptr, nullptr, nullptr, nullptr,
                                 ctx->ev[EV_SORTED], h_num_pairs, &ok));
    if (ok) {
      ctx->tm.bucket_join = 1;
      HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_JOINED], ctx->stream));
      return HMJ_OK;
    }
  }
  // skewed hashes: stable sort of both pair relations, then the merge join
  hmj_relation R = {}, S = {};
  uint64_t* h = nullptr;
  uint32_t* rid = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_R_HASH, nr, &h));
  HMJ_TRY(hmj_ws(ctx, WS_R_ROWID, nr, &rid));
  HMJ_TRY(hmj_impl_sort(ctx, d_rhash, d_rrowid, nr, h, rid));
  R.d_hash = h; R.d_rowid = rid; R.n = nr; R.key_kind = HMJ_KEY_U64;
  HMJ_TRY(hmj_ws(ctx, WS_S_HASH, ns, &h));
  HMJ_TRY(hmj_ws(ctx, WS_S_ROWID, ns, &rid));
  HMJ_TRY(hmj_impl_sort(ctx, d_shash, d_srowid, ns, h, rid));
  S.d_hash = h; S.d_rowid = rid; S.n = ns; S.key_kind = HMJ_KEY_U64;
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_SORTED], ctx->stream));
  HMJ_TRY(join_sorted(ctx, &R, &S, h_num_pairs));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_JOINED], ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_join_strings(hmj_ctx* ctx, const uint64_t* d_roffsets, const uint8_t* d_rchars,
                                size_t nr, const uint64_t* d_soffsets, const uint8_t* d_schars,
                                size_t ns, uint64_t* h_num_pairs) {
  return hmj_join_strings_range(ctx, d_roffsets, d_rchars, nr, d_soffsets, d_schars, ns, 0, h_num_pairs);
}

extern "C" int hmj_join_strings_range(hmj_ctx* ctx, const uint64_t* d_roffsets, const uint8_t* d_rchars,
                                      size_t nr, const uint64_t* d_soffsets, const uint8_t* d_schars,
                                      size_t ns, int prefix_bits, uint64_t* h_num_pairs) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, h_num_pairs != nullptr && prefix_bits >= 0 && prefix_bits <= 16);
  HMJ_ARG(ctx, (nr == 0 || d_roffsets) && (ns == 0 || d_soffsets));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->tm.sort_passes = 0;
  ctx->tm.sort_finishes = 0;
  ctx->tm.sort_fallbacks = 0;
  ctx->tm.sort_uniform = 0;
  ctx->tm.bucket_join = 0;
  ctx->tm.scatter_levels = 0;
  uint64_t *rh = nullptr, *sh = nullptr;
  uint32_t *rr = nullptr, *sr = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_R_HASH, nr, &rh));
  HMJ_TRY(hmj_ws(ctx, WS_R_ROWID, nr, &rr));
  HMJ_TRY(hmj_ws(ctx, WS_S_HASH, ns, &sh));
  HMJ_TRY(hmj_ws(ctx, WS_S_ROWID, ns, &sr));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_START], ctx->stream));
  uint64_t *rt = nullptr, *st = nullptr;  // unsorted hashes
  HMJ_TRY(hmj_ws(ctx, WS_HASH_TMP0, nr, &rt));
  HMJ_TRY(hmj_ws(ctx, WS_HASH_TMP1, ns, &st));
  HMJ_TRY(hmj_hash_strings(ctx, d_roffsets, d_rchars, nr, rt));
  HMJ_TRY(hmj_hash_strings(ctx, d_soffsets, d_schars, ns, st));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_HASHED], ctx->stream));
  {
    int ok = 0;
    HMJ_TRY(hmj_join_uniform_try(ctx, HMJ_KEY_STRING, rt, nullptr, nr, st, nullptr, ns, prefix_bits, d_roffsets, d_rchars,
                                 d_soffsets, d_schars, ctx->ev[EV_SORTED], h_num_pairs, &ok));
    if (ok) {
      ctx->tm.bucket_join = 1;
      HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_JOINED], ctx->stream));
      return HMJ_OK;
    }
  }
  const bool skew_r = ctx->sample_skewed[0], skew_s = ctx->sample_skewed[1];
  ctx->next_sort_skewed = skew_r;
  HMJ_TRY(hmj_sort_pairs(ctx, rt, nullptr, nr, rh, rr));
  ctx->next_sort_skewed = skew_s;
  HMJ_TRY(hmj_sort_pairs(ctx, st, nullptr, ns, sh, sr));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_SORTED], ctx->stream));
  hmj_relation R = {rh, rr, nr, HMJ_KEY_STRING, d_roffsets, d_rchars};
  hmj_relation S = {sh, sr, ns, HMJ_KEY_STRING, d_soffsets, d_schars};
  HMJ_TRY(join_sorted(ctx, &R, &S, h_num_pairs));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_JOINED], ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_join_result(hmj_ctx* ctx, const uint32_t** d_out_r, const uint32_t** d_out_s,
                               uint64_t* num_pairs) {
  if (!ctx) return HMJ_EARG;
  if (d_out_r) *d_out_r = reinterpret_cast<const uint32_t*>(ctx->ws[WS_OUT_R].ptr);
  if (d_out_s) *d_out_s = reinterpret_cast<const uint32_t*>(ctx->ws[WS_OUT_S].ptr);
  if (num_pairs) *num_pairs = ctx->last_pairs;
  return HMJ_OK;
}

extern "C" int hmj_join_result_to_host(hmj_ctx* ctx, uint32_t* h_out_r, uint32_t* h_out_s,
                                       size_t cap) {
  if (!ctx) return HMJ_EARG;
  const size_t m = (size_t)ctx->last_pairs;
  if (m > cap) return hmj_fail(ctx, HMJ_ECAPACITY, "output capacity too small", __FILE__, __LINE__);
  if (m == 0) return HMJ_OK;
  HMJ_ARG(ctx, h_out_r && h_out_s);
  HMJ_CUDA(ctx, cudaMemcpyAsync(h_out_r, ctx->ws[WS_OUT_R].ptr, m * 4, cudaMemcpyDeviceToHost, ctx->stream));
  HMJ_CUDA(ctx, cudaMemcpyAsync(h_out_s, ctx->ws[WS_OUT_S].ptr, m * 4, cudaMemcpyDeviceToHost, ctx->stream));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_join_result_copy(hmj_ctx* ctx, uint32_t* d_out_r, uint32_t* d_out_s, size_t cap) {
  if (!ctx) return HMJ_EARG;
  const size_t m = (size_t)ctx->last_pairs;
  if (m > cap) return hmj_fail(ctx, HMJ_ECAPACITY, "output capacity too small", __FILE__, __LINE__);
  if (m == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_out_r && d_out_s);
  HMJ_CUDA(ctx, cudaMemcpyAsync(d_out_r, ctx->ws[WS_OUT_R].ptr, m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  HMJ_CUDA(ctx, cudaMemcpyAsync(d_out_s, ctx->ws[WS_OUT_S].ptr, m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  return HMJ_OK;
}

extern "C" int hmj_set_join_order(hmj_ctx* ctx, int hash_order) {
  if (!ctx) return HMJ_EARG;
  ctx->hash_order = hash_order != 0;
  return HMJ_OK;
}

extern "C" int hmj_get_timings(hmj_ctx* ctx, hmj_timings* out) {
  if (!ctx || !out) return HMJ_EARG;
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  hmj_timings t = ctx->tm;
  float ms = 0;
  if (cudaEventElapsedTime(&ms, ctx->ev[EV_START], ctx->ev[EV_HASHED]) == cudaSuccess) t.hash_ms = ms;
  if (cudaEventElapsedTime(&ms, ctx->ev[EV_HASHED], ctx->ev[EV_SORTED]) == cudaSuccess) t.sort_ms = ms;
  if (cudaEventElapsedTime(&ms, ctx->ev[EV_SORTED], ctx->ev[EV_JOINED]) == cudaSuccess) t.join_ms = ms;
  if (cudaEventElapsedTime(&ms, ctx->ev[EV_START], ctx->ev[EV_JOINED]) == cudaSuccess) t.total_ms = ms;
  t.pass_ms_sum = 0;
  t.pass_launches = 0;
  for (int k = 0; k < 8; ++k) {
    t.kernel_launches[k] = 0;
    t.kernel_ms[k] = 0;
    t.kernel_bytes[k] = 0;
  }
  for (size_t i = 0; i + 1 < ctx->prof_used; i += 2) {
    if (cudaEventElapsedTime(&ms, ctx->prof_events[i], ctx->prof_events[i + 1]) == cudaSuccess) {
      const int k = ctx->prof_kind[i / 2] & 7;
      t.kernel_launches[k]++;
      t.kernel_ms[k] += ms;
      t.kernel_bytes[k] += ctx->prof_bytes[i / 2];
    }
  }
  t.pass_ms_sum = t.kernel_ms[HMJ_K_PASS];
  t.pass_launches = t.kernel_launches[HMJ_K_PASS];
  cudaGetLastError();  // events never recorded yet: not an error for the caller
  *out = t;
  return HMJ_OK;
}

extern "C" int hmj_set_profiling(hmj_ctx* ctx, int on) {
  if (!ctx) return HMJ_EARG;
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->prof_on = on != 0;
  ctx->prof_used = 0;
  return HMJ_OK;
}
