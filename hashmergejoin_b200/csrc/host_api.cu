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

namespace {

int ensure_copy_streams(hmj_ctx* ctx, size_t events) {
  if (!ctx->copy_in) HMJ_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking));
  if (!ctx->copy_out) HMJ_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking));
  while (ctx->copy_events.size() < events) {
    cudaEvent_t e;
    HMJ_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ctx->copy_events.push_back(e);
  }
  return HMJ_OK;
}

// Chunks of R per host-buffer join.  More chunks shorten the tail (the last chunk's scatter, join
// and copy out, which nothing overlaps) and cost a host round trip each: at 256M rows 4 / 8 / 12 /
// 16 / 24 / 32 chunks take 94.3 / 90.0 / 89.0 / 88.6 / 88.7 / 90.1 ms (profiles/r2z6_e2e_chunks.txt),
// so 16, with chunks of at least 512K rows.  HMJ_HOST_CHUNKS overrides; read per call so that one
// process can sweep it (scripts/e2e_chunks.py).
int host_chunks_for(size_t nr) {
  const char* e = getenv("HMJ_HOST_CHUNKS");
  if (e) {
    const int c = atoi(e);
    return c < 1 ? 1 : c > 32 ? 32 : c;
  }
  const size_t c = nr >> 19;
  return c < 2 ? 2 : c > 16 ? 16 : (int)c;
}

// One host-buffer join with the three legs overlapped.  A join's pairs depend on all of S but
// only on the rows of R seen so far, so: S is copied in and cut into its buckets; R follows in
// chunks, and as soon as a chunk has landed it is cut into the same buckets and joined against
// S's, its pairs appended to the result and copied out while the next chunk is still arriving
// (the host link is full duplex).  The call then costs about the input copy plus one chunk's
// tail instead of copy + join + copy.  The union of the chunks' pairs is the join (every R row
// is in exactly one chunk).  *done = 0: the plan does not apply (small or skewed input, result
// larger than expected) and nothing has been delivered -- the caller runs the plain path, the
// inputs are complete in d_r / d_s by then.
int join_u64_host_chunked(hmj_ctx* ctx, const uint64_t* h_rkeys, size_t nr, const uint64_t* h_skeys,
                          size_t ns, uint32_t* h_out_r, uint32_t* h_out_s, size_t cap,
                          uint64_t* h_num_pairs, uint64_t* d_r, uint64_t* d_s, int* done) {
  *done = 0;
  const int chunks = host_chunks_for(nr);
  hmj_uniform_plan p = hmj_uniform_plan_for(nr > ns ? nr : ns, 0);
  const char* e = getenv("HMJ_JOIN_UNIFORM");
  const bool eligible = chunks > 1 && p.levels != 0 && nr >= ((size_t)chunks << 16) && nr >= (size_t(1) << 19) && ns > 0 &&
                        !(e && atoi(e) == 0) && !ctx->hash_order;
  HMJ_TRY(ensure_copy_streams(ctx, (size_t)chunks + 2));
  cudaStream_t st = ctx->stream, cin = ctx->copy_in, cout = ctx->copy_out;
  // chunk boundaries on multiples of 4096 rows: tiles of the scatter stay 16-byte aligned
  const size_t per = eligible ? ((nr / chunks + 4095) & ~size_t(4095)) : nr;
  // the copy stream must not overtake work still queued on the compute stream that reads the
  // staging buffers of an earlier call
  HMJ_CUDA(ctx, cudaEventRecord(ctx->copy_events[chunks + 1], st));
  HMJ_CUDA(ctx, cudaStreamWaitEvent(cin, ctx->copy_events[chunks + 1], 0));
  if (ns) HMJ_CUDA(ctx, cudaMemcpyAsync(d_s, h_skeys, ns * 8, cudaMemcpyHostToDevice, cin));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->copy_events[chunks], cin));
  int n_chunks = 0;
  for (size_t off = 0; off < nr; off += per, ++n_chunks) {
    const size_t n = nr - off < per ? nr - off : per;
    HMJ_CUDA(ctx, cudaMemcpyAsync(d_r + off, h_rkeys + off, n * 8, cudaMemcpyHostToDevice, cin));
    HMJ_CUDA(ctx, cudaEventRecord(ctx->copy_events[n_chunks], cin));
  }
  if (!eligible) {
    HMJ_CUDA(ctx, cudaStreamSynchronize(cin));
    return HMJ_OK;
  }
  ctx->tm.sort_passes = ctx->tm.sort_finishes = ctx->tm.sort_fallbacks = ctx->tm.sort_uniform = 0;
  ctx->tm.bucket_join = 0;
  ctx->tm.scatter_levels = 0;
  // result buffers and the pair counter
  size_t dcap = ctx->ws[WS_OUT_R].bytes / 4;
  const size_t guess = nr > ns ? nr : ns;
  if (dcap < guess) dcap = guess;
  uint32_t *out_r = nullptr, *out_s = nullptr;
  uint64_t* d_total = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_OUT_R, dcap, &out_r));
  HMJ_TRY(hmj_ws(ctx, WS_OUT_S, dcap, &out_s));
  HMJ_TRY(hmj_ws(ctx, WS_JOIN_TILES, 4, &d_total));
  dcap = ctx->ws[WS_OUT_R].bytes / 4;
  if (ctx->ws[WS_OUT_S].bytes / 4 < dcap) dcap = ctx->ws[WS_OUT_S].bytes / 4;
  HMJ_TRY(hmj_fill_async(ctx, d_total, 0, 32, st));
  // The plan comes from a sample of S and of R's FIRST chunk, taken as soon as both have landed
  // (the other chunks keep arriving meanwhile): a hash prefix all sampled keys share is skipped
  // -- dense integer keys -- and checked on every row by level 1 of every chunk; a relation the
  // sample finds skewed ends the attempt before anything is scattered.
  const size_t n0 = nr < per ? nr : per;
  if (nr + ns >= ((size_t)1 << 22) && n0 >= 32768 && ns >= 32768) {
    HMJ_CUDA(ctx, cudaEventSynchronize(ctx->copy_events[0]));
    int skewed = 0;
    HMJ_TRY(hmj_sample_plan(ctx, d_r, n0, d_s, ns, &p, &skewed));
    p = skewed ? hmj_uniform_plan() : p;
    if (p.levels != 0 && (nr > ns ? nr : ns) != (n0 > ns ? n0 : ns)) {
      // (the sample planned for chunk 0's size: keep its prefix, take the bucket count of the join)
      const double dens = p.density > 1.0 ? p.density : 1.0;
      double want = (double)(nr > ns ? nr : ns) * dens;
      if (dens > 1.0 && want > ldexp(1.0, 64 - p.skip)) want = ldexp(1.0, 64 - p.skip);  // one row per value at most
      hmj_uniform_plan q = hmj_uniform_plan_for((size_t)ceil(want), p.skip);
      q.check_prefix = p.check_prefix;
      q.prefix = p.prefix;
      q.density = p.density;
      p = q;
    }
    if (p.levels == 0) {
      HMJ_CUDA(ctx, cudaStreamSynchronize(cin));
      return HMJ_OK;
    }
  }
  // S into its buckets
  hmj_uniform_buckets Sb, Rb;
  HMJ_CUDA(ctx, cudaStreamWaitEvent(st, ctx->copy_events[chunks], 0));
  HMJ_TRY(hmj_uniform_scatter(ctx, d_s, nullptr, ns, 0, p, 1, &Sb));
  uint64_t* h = reinterpret_cast<uint64_t*>(ctx->pinned);
  uint64_t m_prev = 0;
  bool failed = false;
  int c = 0;
  for (size_t off = 0; off < nr && !failed; off += per, ++c) {
    const size_t n = nr - off < per ? nr - off : per;
    HMJ_CUDA(ctx, cudaStreamWaitEvent(st, ctx->copy_events[c], 0));
    HMJ_TRY(hmj_uniform_scatter(ctx, d_r + off, nullptr, n, (uint32_t)off, p, 0, &Rb));
    HMJ_TRY(hmj_bucket_join_launch(ctx, HMJ_KEY_U64, Rb, Sb, p.cap_seg, 64 - p.skip - p.qb - 12, n, ns,
                                   nullptr, nullptr, nullptr, nullptr, out_r, out_s, dcap, d_total));
    HMJ_CUDA(ctx, cudaMemcpyAsync(h, d_total, 8, cudaMemcpyDeviceToHost, st));
    HMJ_CUDA(ctx, cudaMemcpyAsync(reinterpret_cast<uint32_t*>(h + 1), Rb.flag, 4, cudaMemcpyDeviceToHost, st));
    HMJ_CUDA(ctx, cudaMemcpyAsync(reinterpret_cast<uint32_t*>(h + 1) + 1, Sb.flag, 4, cudaMemcpyDeviceToHost, st));
    HMJ_CUDA(ctx, cudaStreamSynchronize(st));
    const uint64_t m = h[0];
    if (h[1] != 0 || m > dcap || ((h_out_r || h_out_s) && m > cap)) {
      failed = true;  // skewed hashes, or more pairs than the buffers hold: the plain path decides
      break;
    }
    if (m > m_prev) {
      // the chunk's pairs leave while the next chunk arrives (the kernels that wrote them are done)
      if (h_out_r) HMJ_CUDA(ctx, cudaMemcpyAsync(h_out_r + m_prev, out_r + m_prev, (m - m_prev) * 4, cudaMemcpyDeviceToHost, cout));
      if (h_out_s) HMJ_CUDA(ctx, cudaMemcpyAsync(h_out_s + m_prev, out_s + m_prev, (m - m_prev) * 4, cudaMemcpyDeviceToHost, cout));
    }
    m_prev = m;
  }
  HMJ_CUDA(ctx, cudaStreamSynchronize(cin));
  HMJ_CUDA(ctx, cudaStreamSynchronize(cout));
  if (failed) return HMJ_OK;
  ctx->tm.bucket_join = 1;
  ctx->last_pairs = m_prev;
  *h_num_pairs = m_prev;
  *done = 1;
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
  HMJ_TRY(hmj_ws(ctx, WS_IN_0, nr, &dr));
  HMJ_TRY(hmj_ws(ctx, WS_IN_1, ns, &ds));
  {
    // one call at a time per GPU owns the host link's input direction (see CopyTurn)
    CopyTurn turn(ctx, TURN_H2D);
    mark(ctx, "h2d turn");
    int done = 0;
    const int rc = join_u64_host_chunked(ctx, h_rkeys, nr, h_skeys, ns, h_out_r, h_out_s, cap,
                                         h_num_pairs, dr, ds, &done);
    const int rc2 = turn.finish();
    HMJ_TRY(rc);
    HMJ_TRY(rc2);
    if (done) {
      mark(ctx, "leave (chunked)");
      return HMJ_OK;
    }
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

extern "C" int hmj_host_alloc(void** out, size_t bytes) {
  if (!out) return HMJ_EARG;
  *out = nullptr;
  const cudaError_t e = cudaMallocHost(out, bytes ? bytes : 1);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? HMJ_ENOMEM : HMJ_ECUDA;
  }
  return HMJ_OK;
}

extern "C" int hmj_host_free(void* p) {
  if (p && cudaFreeHost(p) != cudaSuccess) {
    cudaGetLastError();
    return HMJ_ECUDA;
  }
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
