// common.cuh -- context, error handling and workspace for libhmj.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/hmj.h"

#define HMJ_NUM_SMS_HINT 148

// Named scratch slots of the context workspace (grow-only, reused across calls so that no
// cudaMalloc happens in a steady-state timed region).
enum hmj_ws_slot {
  WS_HIST = 0,      // digit histograms / bin offsets
  WS_LOOKBACK,      // decoupled look-back tile descriptors
  WS_SORT_KEYS_A,   // ping-pong buffers of the sort
  WS_SORT_VALS_A,
  WS_SORT_KEYS_B,
  WS_SORT_VALS_B,
  WS_R_HASH,        // whole-operator: sorted columns of R and S
  WS_R_ROWID,
  WS_S_HASH,
  WS_S_ROWID,
  WS_JOIN_TILES,    // merge-path partition points + per-tile counts
  WS_JOIN_MISC,
  WS_OUT_R,         // whole-operator: result pairs
  WS_OUT_S,
  WS_IN_0,          // host-entry staging of inputs
  WS_IN_1,
  WS_IN_2,
  WS_IN_3,
  WS_TABLE,         // hash-join baseline table
  WS_SCALARS,       // small device scalars (counters, sums)
  WS_SEG_WINDOWS,   // window table of the segment sort
  WS_US_CURSORS,    // uniform sort: flag, bucket cursors, dense offsets
  WS_US_CURSORS_1,  // ... of the second relation (bucket join)
  WS_SORT_KEYS_C,   // padded buckets of the second relation (bucket join)
  WS_SORT_VALS_C,
  WS_HASH_TMP0,     // unsorted hash columns (strings), so that the sort never runs in place
  WS_HASH_TMP1,
  WS_NUM_SLOTS
};

struct hmj_buf {
  void* ptr = nullptr;
  size_t bytes = 0;
};

struct hmj_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int num_sms = HMJ_NUM_SMS_HINT;
  std::string err;
  hmj_buf ws[WS_NUM_SLOTS];
  void* pinned = nullptr;  // small pinned host scratch for scalar read-backs
  size_t pinned_bytes = 0;
  uint64_t launches = 0;
  // decoupled look-back epoch (radix_sort.cu): descriptors carry it so no per-pass memset
  uint32_t lb_epoch = 0;
  void* lb_seen_ptr = nullptr;
  size_t lb_seen_bytes = 0;
  // optional per-launch timing of the dominant kernel (onesweep pass), see hmj_set_profiling
  bool prof_on = false;
  std::vector<cudaEvent_t> prof_events;  // pairs: [2i] before, [2i+1] after launch i
  std::vector<int> prof_kind;            // kernel class of pair i (hmj_kernel_class)
  std::vector<double> prof_bytes;        // algorithmic bytes of launch i
  size_t prof_used = 0;
  // hmj_set_join_order: pairs wanted in ascending hash order (the reference's iteration order,
  // hashjoin.h:126-153): joins then take the sort + merge-path plan
  bool hash_order = false;
  // set by a join whose sample found a relation skewed, consumed by the next sort of pairs
  bool next_sort_skewed = false;
  bool sample_skewed[2] = {false, false};  // (R, S) verdict of the last hmj_sample_plan
  std::vector<uint64_t> sample_host;  // page-locked landing area of hmj_sample_plan
  // copy streams and events of the host-buffer entry points (host_api.cu), created on first use
  cudaStream_t copy_in = nullptr, copy_out = nullptr;
  std::vector<cudaEvent_t> copy_events;
  // multi-GPU join state (dist_join.cu)
  struct hmj_dist_state* dist = nullptr;
  struct hmj_xchg_state* xchg = nullptr;  // string-key exchange over peer memory (dist_strings.cu)
  // last whole-operator result
  uint64_t last_pairs = 0;
  hmj_timings tm;
  cudaEvent_t ev[8] = {};
};

static inline int hmj_fail(hmj_ctx* ctx, int code, const char* what, const char* file, int line) {
  if (ctx) {
    char b[512];
    snprintf(b, sizeof b, "%s (%s:%d)", what, file, line);
    ctx->err = b;
  }
  return code;
}

#define HMJ_CUDA(ctx, call)                                                        \
  do {                                                                             \
    cudaError_t e__ = (call);                                                      \
    if (e__ != cudaSuccess) {                                                      \
      return hmj_fail(ctx, e__ == cudaErrorMemoryAllocation ? HMJ_ENOMEM : HMJ_ECUDA, \
                      cudaGetErrorString(e__), __FILE__, __LINE__);                \
    }                                                                              \
  } while (0)

#define HMJ_TRY(expr)            \
  do {                           \
    int rc__ = (expr);           \
    if (rc__ != HMJ_OK) return rc__; \
  } while (0)

#define HMJ_ARG(ctx, cond)                                                   \
  do {                                                                       \
    if (!(cond)) return hmj_fail(ctx, HMJ_EARG, "bad argument: " #cond, __FILE__, __LINE__); \
  } while (0)

// checks the launch itself (not completion) and counts it
#define HMJ_LAUNCHED(ctx)                 \
  do {                                    \
    (ctx)->launches++;                    \
    HMJ_CUDA(ctx, cudaGetLastError());    \
  } while (0)

// returns a device scratch buffer of at least `bytes` bytes in slot `slot`
int hmj_ws_get(hmj_ctx* ctx, int slot, size_t bytes, void** out);
// Fill device memory with a byte value by a kernel on `st` (bytes a multiple of 4, ptr 4-byte
// aligned): control words are cleared on the SMs, in stream order, whatever the copy engines do.
int hmj_fill_async(hmj_ctx* ctx, void* ptr, int byte, size_t bytes, cudaStream_t st);
// HMJ_HOST_TRACE=1: wall-clock marks of host-side progress on stderr (debugging aid)
void hmj_host_mark(hmj_ctx* ctx, const char* what);

template <typename T>
static inline int hmj_ws(hmj_ctx* ctx, int slot, size_t count, T** out) {
  void* p = nullptr;
  int rc = hmj_ws_get(ctx, slot, count * sizeof(T), &p);
  *out = reinterpret_cast<T*>(p);
  return rc;
}

// Per-launch timing of the hot kernels while hmj_set_profiling is on: bracket a launch with
//   cudaEvent_t after; HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_x, bytes, &after)); <launch>; hmj_prof_end(ctx, after);
// `bytes` = algorithmic bytes the launch moves.
static inline int hmj_prof_begin_on(hmj_ctx* ctx, int kind, double bytes, cudaEvent_t* after,
                                    cudaStream_t st) {
  *after = nullptr;
  if (!ctx->prof_on) return HMJ_OK;
  if (ctx->prof_used + 2 > ctx->prof_events.size()) {
    cudaEvent_t a, b;
    HMJ_CUDA(ctx, cudaEventCreate(&a));
    HMJ_CUDA(ctx, cudaEventCreate(&b));
    ctx->prof_events.push_back(a);
    ctx->prof_events.push_back(b);
    ctx->prof_kind.push_back(0);
    ctx->prof_bytes.push_back(0);
  }
  ctx->prof_kind[ctx->prof_used / 2] = kind;
  ctx->prof_bytes[ctx->prof_used / 2] = bytes;
  HMJ_CUDA(ctx, cudaEventRecord(ctx->prof_events[ctx->prof_used], st));
  *after = ctx->prof_events[ctx->prof_used + 1];
  ctx->prof_used += 2;
  return HMJ_OK;
}
static inline void hmj_prof_end_on(hmj_ctx* ctx, cudaEvent_t after, cudaStream_t st) {
  if (after) cudaEventRecord(after, st);
}
static inline int hmj_prof_begin(hmj_ctx* ctx, int kind, double bytes, cudaEvent_t* after) {
  return hmj_prof_begin_on(ctx, kind, bytes, after, ctx->stream);
}
static inline void hmj_prof_end(hmj_ctx* ctx, cudaEvent_t after) {
  hmj_prof_end_on(ctx, after, ctx->stream);
}

static inline size_t hmj_div_up(size_t a, size_t b) { return (a + b - 1) / b; }

// ---- stage entry points shared between translation units -----------------------------
// (implemented in hash.cu / radix_sort.cu / merge_join.cu / hash_table.cu / partition.cu)
int hmj_impl_sort(hmj_ctx* ctx, const uint64_t* d_keys_in, const uint32_t* d_vals_in, size_t n,
                  uint64_t* d_keys_out, uint32_t* d_vals_out);
// segment_sort.cu
int hmj_segment_sort_cap();
int hmj_launch_segment_sort(hmj_ctx* ctx, uint64_t* keys, uint32_t* vals, uint32_t n, int q,
                            int group_shift, uint32_t stride, uint32_t* d_flag);
int hmj_launch_bucket_finish(hmj_ctx* ctx, uint64_t* keys, uint32_t* vals, const uint32_t* counts,
                             const uint32_t* offsets, uint32_t buckets, uint32_t in_cap, int q,
                             uint64_t* out_keys, uint32_t* out_vals, uint32_t* d_flag);
// bucket_scatter.cu
// regions of a scatter level that live on the GPUs of a multi-GPU job (peer pointers)
struct hmj_peer_regions {
  uint64_t* keys[8];
  uint32_t* vals[8];
  uint32_t* counts[8];  // per owner: pairs in sub-region (region * world + source rank)
  int own_bits;         // log2(regions per GPU)
  int world, rank;      // GPUs of the job, this GPU
  int bulk;             // runs leave as bulk async stores (HMJ_DIST_BULK, default on)
};
int hmj_scatter_level1_dist(hmj_ctx* ctx, cudaStream_t st, const uint64_t* keys_in, size_t n,
                            uint32_t iota_base, int shift, int bits, uint32_t out_cap,
                            const hmj_peer_regions& peers, uint32_t* d_cursors, uint32_t* d_flag,
                            uint32_t* d_tile_counter, uint32_t count_table_offset,
                            uint32_t ctas_per_sm);
int hmj_scatter_level2(hmj_ctx* ctx, cudaStream_t st, const uint64_t* k1, const uint32_t* v1,
                       const uint32_t* cur1, const uint32_t* start1, uint32_t cap1, uint32_t regions,
                       size_t n_hint, int shift, int bits, uint64_t* k2, uint32_t* v2,
                       uint32_t cap_seg, uint32_t* cur2, uint32_t* d_flag, uint32_t* d_tile_counter,
                       int region_shift);
struct hmj_uniform_plan {
  int levels;        // 0 = not eligible
  int bits[2];
  int qb;            // total prefix bits = bits[0] + bits[1]
  int skip;          // topmost hash bits shared by all rows (multi-GPU hash range), not used
  uint32_t cap_seg;  // slots per final bucket
  // skip found by sampling (dense integer keys): level 1 checks every key's top `skip` bits
  // against `prefix` and voids the attempt when one differs
  bool check_prefix;
  uint64_t prefix;
  // > 1 (sample): the keys fill only 1/density of the range under the prefix, evenly -- dense ids
  // 0..N-1 with N not a power of two.  A region then holds n * density / regions rows and the
  // bucket count was chosen for n * density rows.  0 or 1: the rows spread over the whole range.
  double density;
};
struct hmj_uniform_buckets {
  uint64_t* keys;     // bucket b at [b * cap_seg ...)
  uint32_t* vals;
  uint32_t* counts;   // [buckets]
  uint32_t* offsets;  // [buckets + 1] scratch for the dense offsets
  uint32_t* flag;     // != 0: a region overflowed
  uint32_t buckets;
};
hmj_uniform_plan hmj_uniform_plan_for(size_t n_plan, int skip_bits);
int hmj_uniform_scatter(hmj_ctx* ctx, const uint64_t* keys_in, const uint32_t* vals_in, size_t n,
                        uint32_t iota_base, const hmj_uniform_plan& p, int set,
                        hmj_uniform_buckets* out);
// bucket_join.cu
int hmj_join_uniform_try(hmj_ctx* ctx, int kind, const uint64_t* r_hash, const uint32_t* r_rowid,
                         size_t nr, const uint64_t* s_hash, const uint32_t* s_rowid, size_t ns,
                         int skip_bits, const uint64_t* roff,
                         const uint8_t* rchars, const uint64_t* soff, const uint8_t* schars,
                         cudaEvent_t ev_partitioned, uint64_t* h_num_pairs, int* ok);
int hmj_bucket_join_launch(hmj_ctx* ctx, int kind, const hmj_uniform_buckets& R,
                           const hmj_uniform_buckets& S, uint32_t cap_seg, int dshift, size_t nr,
                           size_t ns, const uint64_t* roff, const uint8_t* rchars,
                           const uint64_t* soff, const uint8_t* schars, uint32_t* out_r,
                           uint32_t* out_s, size_t cap, uint64_t* d_total);
int hmj_bucket_join_run(hmj_ctx* ctx, int kind, const hmj_uniform_buckets& R,
                        const hmj_uniform_buckets& S, uint32_t cap_seg, int dshift, size_t nr,
                        size_t ns, const uint64_t* roff, const uint8_t* rchars,
                        const uint64_t* soff, const uint8_t* schars, uint64_t* h_num_pairs,
                        int* ok);
int hmj_sample_plan(hmj_ctx* ctx, const uint64_t* r_hash, size_t nr, const uint64_t* s_hash,
                    size_t ns, hmj_uniform_plan* p, int* skewed);
// the host half of hmj_sample_plan: h = per_side evenly spaced hashes of R, then of S
void hmj_plan_from_sample(const uint64_t* h, uint32_t per_side, size_t nr, size_t ns,
                          hmj_uniform_plan* p, bool skewed_side[2]);
int hmj_sort_uniform_try(hmj_ctx* ctx, const uint64_t* keys_in, size_t n, uint32_t iota_base,
                         uint64_t* keys_out, uint32_t* vals_out, int* ok);
int hmj_impl_join_sorted(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s,
                         uint64_t* h_num_pairs);
