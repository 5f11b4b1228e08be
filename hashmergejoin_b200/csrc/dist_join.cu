// dist_join.cu -- multi-GPU hash-merge join with the exchange FUSED into the first scatter level
// and the two relations PIPELINED against each other.
//
// every GPU splits its rows of R and S by hash range, one exchange moves range g
// to GPU g, then every GPU joins its own range.  Here the split and the exchange are one kernel:
// level 1 of the bucket scatter (bucket_scatter.cu) runs over the GPU's local rows, cuts them on
// the top log2(G) hash bits -- exactly partition_only's h >> (64 - bits), partitioned_hash.h:59 --
// and writes every run straight into the owner's memory through NVLink peer pointers (a 4096-pair
// tile leaves as G runs of ~4096/G pairs: KBs per remote store).  Every source GPU owns a
// sub-region of every destination (slot reservation = atomicAdd on a cursor in the SOURCE's
// memory; the counts are handed to the owners after the kernel).  No partition pass, no send
// buffers, no NCCL all-to-all.  It stands in for partition_only's per-thread push_back into
// shared partitions (partitioned_hash.h:82-124) at the scale of a node.
//
// One join is ONE call (hmj_dist_join_u64) that only enqueues work; the GPUs meet through flags in
// each other's memory, not through host barriers.  Each relation crosses NVLink in PARTS (2 by
// default, HMJ_DIST_PARTS), so that the owners always have a part to cut locally while the next
// one is in flight:
//   exchange stream  [wait DONE(R, last join)] level 1 of R part 0 over NVLink -> counts -> READY(R,0)
//                    level 1 of R part 1 -> counts -> READY(R,1); the same for S
//   main stream      wait READY(R,0) from all -> local levels of what R part 0 brought -> ... R part 1
//                    -> DONE(R) to all -> S part 0 -> S part 1 -> DONE(S) to all -> bucket join
// A part appends to the same sub-regions (the source's cursors run on); the owner keeps one count
// table per part, so its local level of part p reads [count(p-1), count(p)) of every sub-region
// and appends to the final buckets.  READY(side, part) from source s = "my runs and counts of this
// part are in your memory" (release store after the kernels that wrote them; the owner acquires
// it in a one-CTA wait kernel in front of its local level).  DONE(side) from owner o = "I have
// read your runs of this join": the source's next level 1 of that side waits for it before it
// overwrites them.  A wait gives up after HMJ_DIST_TIMEOUT_MS (default 20000) and fails the join
// instead of hanging the GPU.
// All level-1 launches but the first run with one CTA per SM and the scatter kernels take their
// tiles from a counter, so the local levels fill the other half of every SM while a part is in
// flight and all of it afterwards.
//
// One process per GPU: the regions are cudaMalloc'ed here and shared with CUDA IPC handles that
// the caller exchanges over torch.distributed (hashmergejoin_b200/dist.py).
#include <math.h>
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"

namespace {
constexpr int HANDLE_BYTES = 64;  // sizeof(cudaIpcMemHandle_t)
static_assert(sizeof(cudaIpcMemHandle_t) == HANDLE_BYTES, "IPC handle size");
constexpr int HANDLES_PER_RANK = 7;  // per side: keys, row ids, counts; then the signal words
constexpr int MAX_PARTS = 4;
// signal words of a rank, 8 ranks at most: READY [side][part][source rank], DONE [side][owner rank]
inline int sig_ready(int side, int part, int rank) { return (side * MAX_PARTS + part) * 8 + rank; }
inline int sig_done(int side, int rank) { return 2 * MAX_PARTS * 8 + side * 8 + rank; }
constexpr int SIG_WORDS = 2 * MAX_PARTS * 8 + 16;

struct sig_ptrs {
  uint32_t* p[8];
};

// after the kernels that wrote into the peers' memory: tell every GPU of the job
__global__ void dist_signal_kernel(const sig_ptrs sig, int world, int slot, uint32_t epoch) {
  const int r = threadIdx.x;
  if (r >= world) return;
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(sig.p[r] + slot), "r"(epoch) : "memory");
}

// in front of the kernels that read what the peers wrote: all `world` words at `mine` >= epoch
__global__ void dist_wait_kernel(const uint32_t* mine, int world, uint32_t epoch,
                                 uint32_t* fail_flag, unsigned long long timeout_ns) {
  const int r = threadIdx.x;
  if (r >= world) return;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine + r) : "memory");
    if ((int32_t)(v - epoch) >= 0) break;
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) {
      atomicOr(fail_flag, 4u);
      break;
    }
    __nanosleep(100);
  }
}
}  // namespace

struct hmj_dist_state {
  int world = 1, rank = 0;
  // level-1 bits (job wide: the GPU bits + own_bits), local level bits (b3 = 0: one local level)
  int g_bits = 0, b1 = 0, b2 = 0, b3 = 0, own_bits = 0;
  uint32_t cap1 = 0, cap2 = 0, cap_seg = 0, regions_own = 0;
  int parts = 1;          // parts a relation crosses NVLink in (HMJ_DIST_PARTS)
  uint32_t ctab = 0;      // words per count table (one table per part)
  // own level-1 regions per side (0 = R, 1 = S): regions_own x world sub-regions of cap1 slots
  // (keys, row ids) and their counts [regions_own * world], all shared with the peers;
  uint64_t* k1[2] = {nullptr, nullptr};
  uint32_t* v1[2] = {nullptr, nullptr};
  uint32_t* c1[2] = {nullptr, nullptr};
  // private: this source's cursors over all 2^b1 regions of the job (+ 8 words: [2^b1] = flag,
  // [2^b1 + 1 + p] = tile counter of level 1 of part p)
  uint32_t* cur[2] = {nullptr, nullptr};
  uint32_t* sig = nullptr;  // this rank's signal words (shared with the peers)
  sig_ptrs sigs = {};
  hmj_peer_regions peers[2];
  void* opened[8][HANDLES_PER_RANK] = {};
  size_t rows_total = 0;
  uint32_t epoch = 0;
  cudaStream_t side_stream = nullptr;  // the exchange stream
  cudaEvent_t ev_r_sent = nullptr, ev_r_local = nullptr;  // ev_r_sent: fork main -> exchange
  unsigned long long timeout_ns = 20000ull * 1000000ull;
  // what the last join moved (read back from the cursors and the owners' count tables)
  uint64_t sent_pairs = 0, received_pairs[2] = {0, 0};
};

extern "C" int hmj_dist_handle_bytes(void) { return HANDLES_PER_RANK * HANDLE_BYTES; }

// Plans the job for `rows_total` rows per relation over `world` GPUs, allocates this rank's
// level-1 regions and writes their 7 IPC handles (R keys, R row ids, R counts, S ..., signals)
// to handles_out[hmj_dist_handle_bytes()].
extern "C" int hmj_dist_setup(hmj_ctx* ctx, int rank, int world, size_t rows_total,
                              unsigned char* handles_out) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, world >= 2 && world <= 8 && (world & (world - 1)) == 0 && rank >= 0 && rank < world);
  HMJ_ARG(ctx, handles_out != nullptr && ctx->dist == nullptr);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  hmj_dist_state* d = new hmj_dist_state();
  d->world = world;
  d->rank = rank;
  d->rows_total = rows_total;
  while ((1 << d->g_bits) < world) ++d->g_bits;
  int total_bits = 0;
  while (((rows_total + ((size_t)1 << total_bits) - 1) >> total_bits) > 4096) ++total_bits;
  if (total_bits > 18 || total_bits < d->g_bits + 2) {
    delete d;
    return hmj_fail(ctx, HMJ_EARG, "dist join: rows_total outside the supported range", __FILE__, __LINE__);
  }
  // Level 1 splits on the GPU bits only: a tile then leaves as `world` runs of ~tile/world pairs
  // (KBs per remote store instead of the 128 B of a 2^9-way split), and the bits below are cut
  // locally in one or two more levels.  HMJ_DIST_L1_BITS widens level 1 (measurements).
  // Level 1 takes one more bit only where the local levels would otherwise need a 9-bit cut (the
  // 1024-thread configuration of the scatter runs at 0.47 of the HBM peak against 0.68 for the
  // 8-bit one: C4 on 2 GPUs).
  d->b1 = d->g_bits;
  if (total_bits - d->b1 > 16) d->b1 = total_bits - 16;
  if (const char* e = getenv("HMJ_DIST_L1_BITS")) {
    const int v = atoi(e);
    if (v > d->b1 && v <= 9 && v <= total_bits - 1) d->b1 = v;
  }
  if (const char* e = getenv("HMJ_DIST_PARTS")) {
    const int v = atoi(e);
    if (v >= 1 && v <= MAX_PARTS) d->parts = v;
  }
  if (const char* e = getenv("HMJ_DIST_TIMEOUT_MS")) {
    const long v = atol(e);
    if (v > 0) d->timeout_ns = (unsigned long long)v * 1000000ull;
  }
  const int rem = total_bits - d->b1;
  if (rem <= 9) {
    d->b2 = rem;
    d->b3 = 0;
  } else {
    d->b2 = (rem + 1) / 2;
    d->b3 = rem - d->b2;
  }
  d->own_bits = d->b1 - d->g_bits;
  d->regions_own = 1u << d->own_bits;
  if (d->b3 && ((size_t)d->regions_own << d->b2) > 512) {
    delete d;
    return hmj_fail(ctx, HMJ_EARG, "dist join: level-1 width leaves too many local regions", __FILE__, __LINE__);
  }
  if (d->b3) {
    const double avg2 = (double)rows_total / (double)world / (double)((size_t)d->regions_own << d->b2);
    d->cap2 = ((uint32_t)(avg2 + 8.0 * sqrt(avg2) + 64.0) + 255u) & ~255u;
  }
  // a sub-region holds what ONE source sends to a region
  const double avg1 = (double)rows_total / (double)world / (double)(1u << d->b1);
  d->cap1 = ((uint32_t)(avg1 + 8.0 * sqrt(avg1) + 64.0) + 255u) & ~255u;
  d->cap_seg = (uint32_t)hmj_segment_sort_cap();
  ctx->dist = d;  // from here on hmj_dist_teardown frees what has been allocated
  HMJ_CUDA(ctx, cudaStreamCreateWithFlags(&d->side_stream, cudaStreamNonBlocking));
  HMJ_CUDA(ctx, cudaEventCreateWithFlags(&d->ev_r_sent, cudaEventDisableTiming));
  HMJ_CUDA(ctx, cudaEventCreateWithFlags(&d->ev_r_local, cudaEventDisableTiming));
  for (int s = 0; s < 2; ++s) {
    const size_t subs = (size_t)d->regions_own * world;
    const size_t slots = subs * d->cap1;
    HMJ_CUDA(ctx, cudaMalloc(&d->k1[s], slots * 8));
    HMJ_CUDA(ctx, cudaMalloc(&d->v1[s], slots * 4));
    d->ctab = (uint32_t)((subs + 7) & ~size_t(7));
    HMJ_CUDA(ctx, cudaMalloc(&d->c1[s], (size_t)MAX_PARTS * d->ctab * 4));
    HMJ_CUDA(ctx, cudaMemset(d->c1[s], 0, (size_t)MAX_PARTS * d->ctab * 4));
    HMJ_CUDA(ctx, cudaMalloc(&d->cur[s], (((size_t)1 << d->b1) + 8) * 4));
    void* ptrs[3] = {d->k1[s], d->v1[s], d->c1[s]};
    for (int j = 0; j < 3; ++j) {
      cudaIpcMemHandle_t h;
      HMJ_CUDA(ctx, cudaIpcGetMemHandle(&h, ptrs[j]));
      memcpy(handles_out + (size_t)(s * 3 + j) * HANDLE_BYTES, &h, HANDLE_BYTES);
    }
  }
  HMJ_CUDA(ctx, cudaMalloc(&d->sig, SIG_WORDS * 4));
  HMJ_CUDA(ctx, cudaMemset(d->sig, 0, SIG_WORDS * 4));
  {
    cudaIpcMemHandle_t h;
    HMJ_CUDA(ctx, cudaIpcGetMemHandle(&h, d->sig));
    memcpy(handles_out + (size_t)6 * HANDLE_BYTES, &h, HANDLE_BYTES);
  }
  HMJ_CUDA(ctx, cudaDeviceSynchronize());  // the memsets have landed before any peer writes here
  return HMJ_OK;
}

// all_handles[world][7][64]: what every rank wrote in hmj_dist_setup, in rank order.
extern "C" int hmj_dist_connect(hmj_ctx* ctx, const unsigned char* all_handles) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->dist != nullptr && all_handles != nullptr);
  hmj_dist_state* d = ctx->dist;
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  for (int s = 0; s < 2; ++s) {
    memset(&d->peers[s], 0, sizeof d->peers[s]);
    d->peers[s].own_bits = d->own_bits;
    d->peers[s].world = d->world;
    d->peers[s].rank = d->rank;
    const char* e = getenv("HMJ_DIST_BULK");
    d->peers[s].bulk = !(e && atoi(e) == 0);
  }
  for (int r = 0; r < d->world; ++r) {
    void* ptrs[HANDLES_PER_RANK];
    if (r == d->rank) {
      for (int s = 0; s < 2; ++s) {
        ptrs[s * 3 + 0] = d->k1[s]; ptrs[s * 3 + 1] = d->v1[s]; ptrs[s * 3 + 2] = d->c1[s];
      }
      ptrs[6] = d->sig;
    } else {
      for (int j = 0; j < HANDLES_PER_RANK; ++j) {
        cudaIpcMemHandle_t h;
        memcpy(&h, all_handles + ((size_t)r * HANDLES_PER_RANK + j) * HANDLE_BYTES, HANDLE_BYTES);
        HMJ_CUDA(ctx, cudaIpcOpenMemHandle(&ptrs[j], h, cudaIpcMemLazyEnablePeerAccess));
        d->opened[r][j] = ptrs[j];
      }
    }
    for (int s = 0; s < 2; ++s) {
      d->peers[s].keys[r] = reinterpret_cast<uint64_t*>(ptrs[s * 3 + 0]);
      d->peers[s].vals[r] = reinterpret_cast<uint32_t*>(ptrs[s * 3 + 1]);
      d->peers[s].counts[r] = reinterpret_cast<uint32_t*>(ptrs[s * 3 + 2]);
    }
    d->sigs.p[r] = reinterpret_cast<uint32_t*>(ptrs[6]);
  }
  return HMJ_OK;
}

namespace {

int dist_signal(hmj_ctx* ctx, cudaStream_t st, int slot) {
  hmj_dist_state* d = ctx->dist;
  dist_signal_kernel<<<1, 32, 0, st>>>(d->sigs, d->world, slot, d->epoch);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

// all `world` words from d->sig + slot0 on must reach `epoch`
int dist_wait(hmj_ctx* ctx, cudaStream_t st, int slot0, int side, uint32_t epoch) {
  hmj_dist_state* d = ctx->dist;
  uint32_t* fail = d->cur[side] + ((size_t)1 << d->b1);
  dist_wait_kernel<<<1, 32, 0, st>>>(d->sig + slot0, d->world, epoch, fail, d->timeout_ns);
  HMJ_LAUNCHED(ctx);
  return HMJ_OK;
}

// local levels of what part `part` of `side` brought into this rank's sub-regions -> appended to
// the padded final buckets.  words: [8 tile counters][final cursors: buckets][middle: regions2]
int dist_local_levels(hmj_ctx* ctx, cudaStream_t st, int s, int part, hmj_uniform_buckets* out) {
  hmj_dist_state* d = ctx->dist;
  const uint32_t regions2 = d->regions_own << d->b2;  // after the first local level
  const uint32_t buckets = regions2 << d->b3;         // final
  const size_t n_hint = d->rows_total / d->world / d->parts;
  uint32_t* words = reinterpret_cast<uint32_t*>(ctx->ws[s ? WS_US_CURSORS_1 : WS_US_CURSORS].ptr);
  uint64_t* k2 = reinterpret_cast<uint64_t*>(ctx->ws[s ? WS_SORT_KEYS_C : WS_SORT_KEYS_B].ptr);
  uint32_t* v2 = reinterpret_cast<uint32_t*>(ctx->ws[s ? WS_SORT_VALS_C : WS_SORT_VALS_B].ptr);
  uint32_t* flag = d->cur[s] + ((size_t)1 << d->b1);  // this rank's level-1 overflows land here too
  const uint32_t subs = d->regions_own * (uint32_t)d->world;
  const uint32_t* ends = d->c1[s] + (size_t)part * d->ctab;
  const uint32_t* starts = part ? d->c1[s] + (size_t)(part - 1) * d->ctab : nullptr;
  const bool last = part == d->parts - 1;
  if (d->b3 == 0) {
    HMJ_TRY(hmj_scatter_level2(ctx, st, d->k1[s], d->v1[s], ends, starts, d->cap1, subs, n_hint,
                               64 - d->b1 - d->b2, d->b2, k2, v2, d->cap_seg, words + 8, flag,
                               words + 2 * part, d->g_bits));
    if (last) HMJ_TRY(dist_signal(ctx, st, sig_done(s, d->rank)));
  } else {
    // sub-regions -> regions2 middle regions (one buffer, refilled by every part) -> final buckets
    uint64_t* km = reinterpret_cast<uint64_t*>(ctx->ws[WS_SORT_KEYS_A].ptr);
    uint32_t* vm = reinterpret_cast<uint32_t*>(ctx->ws[WS_SORT_VALS_A].ptr);
    uint32_t* cur_mid = words + 8 + buckets;
    HMJ_TRY(hmj_fill_async(ctx, cur_mid, 0, (size_t)regions2 * 4, st));
    HMJ_TRY(hmj_scatter_level2(ctx, st, d->k1[s], d->v1[s], ends, starts, d->cap1, subs, n_hint,
                               64 - d->b1 - d->b2, d->b2, km, vm, d->cap2, cur_mid, flag,
                               words + 2 * part, d->g_bits));
    if (last) HMJ_TRY(dist_signal(ctx, st, sig_done(s, d->rank)));
    HMJ_TRY(hmj_scatter_level2(ctx, st, km, vm, cur_mid, nullptr, d->cap2, regions2, n_hint,
                               64 - d->b1 - d->b2 - d->b3, d->b3, k2, v2, d->cap_seg, words + 8, flag,
                               words + 2 * part + 1, 0));
  }
  out->keys = k2;
  out->vals = v2;
  out->counts = words + 8;
  out->offsets = nullptr;
  out->flag = flag;
  out->buckets = buckets;
  return HMJ_OK;
}

}  // namespace

// The whole multi-GPU join of this rank's blocks of R and S (global row ids r_rowid_base + i and
// s_rowid_base + i).  Every rank of the job calls it once per join, in the same order; nothing
// else synchronises the ranks.  *ok = 0: a region or bucket overflowed on this rank, or a peer did
// not answer in time (the caller agrees on that across ranks and takes the NCCL path).  On
// success this rank's pairs (its hash range) are in the context's result buffers.
extern "C" int hmj_dist_join_u64(hmj_ctx* ctx, const uint64_t* d_rkeys, size_t nr,
                                 uint32_t r_rowid_base, const uint64_t* d_skeys, size_t ns,
                                 uint32_t s_rowid_base, uint64_t* h_num_pairs, int* ok) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->dist != nullptr && h_num_pairs && ok);
  HMJ_ARG(ctx, (nr == 0 || d_rkeys) && (ns == 0 || d_skeys));
  HMJ_ARG(ctx, ((reinterpret_cast<uintptr_t>(d_rkeys) | reinterpret_cast<uintptr_t>(d_skeys)) & 15) == 0);
  hmj_dist_state* d = ctx->dist;
  HMJ_ARG(ctx, d->sigs.p[d->rank] != nullptr);  // hmj_dist_connect has run
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream, sx = d->side_stream;
  *ok = 0;
  *h_num_pairs = 0;
  ctx->tm.sort_passes = ctx->tm.sort_finishes = ctx->tm.sort_fallbacks = ctx->tm.sort_uniform = 0;
  ctx->tm.bucket_join = 0;
  ctx->tm.scatter_levels = 0;

  // every buffer of the join, before anything is enqueued (growing one synchronises the stream)
  const uint32_t regions2 = d->regions_own << d->b2, buckets = regions2 << d->b3;
  const size_t nwords = (size_t)8 + buckets + (d->b3 ? regions2 : 0);
  for (int s = 0; s < 2; ++s) {
    void* p = nullptr;
    HMJ_TRY(hmj_ws_get(ctx, s ? WS_US_CURSORS_1 : WS_US_CURSORS, nwords * 4, &p));
    HMJ_TRY(hmj_ws_get(ctx, s ? WS_SORT_KEYS_C : WS_SORT_KEYS_B, (size_t)buckets * d->cap_seg * 8, &p));
    HMJ_TRY(hmj_ws_get(ctx, s ? WS_SORT_VALS_C : WS_SORT_VALS_B, (size_t)buckets * d->cap_seg * 4, &p));
  }
  if (d->b3) {
    void* p = nullptr;
    HMJ_TRY(hmj_ws_get(ctx, WS_SORT_KEYS_A, (size_t)regions2 * d->cap2 * 8, &p));
    HMJ_TRY(hmj_ws_get(ctx, WS_SORT_VALS_A, (size_t)regions2 * d->cap2 * 4, &p));
  }

  const uint32_t epoch = ++d->epoch;
  const size_t cur_words = ((size_t)1 << d->b1) + 8;
  enum { EV_START = 0, EV_HASHED, EV_SORTED, EV_JOINED };
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_START], st));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_HASHED], st));  // identity hash: no kernel
  for (int s = 0; s < 2; ++s) {
    HMJ_TRY(hmj_fill_async(ctx, d->cur[s], 0, cur_words * 4, st));
    HMJ_TRY(hmj_fill_async(ctx, ctx->ws[s ? WS_US_CURSORS_1 : WS_US_CURSORS].ptr, 0, nwords * 4, st));
  }
  HMJ_CUDA(ctx, cudaEventRecord(d->ev_r_sent, st));
  HMJ_CUDA(ctx, cudaStreamWaitEvent(sx, d->ev_r_sent, 0));

  // exchange stream: both relations, part by part, over NVLink
  const char* e1 = getenv("HMJ_DIST_S_CTAS");
  const uint32_t later_ctas = e1 ? (uint32_t)atoi(e1) : 1u;
  for (int s = 0; s < 2; ++s) {
    const uint64_t* keys = s ? d_skeys : d_rkeys;
    const size_t n = s ? ns : nr;
    const uint32_t base = s ? s_rowid_base : r_rowid_base;
    uint32_t* flag = d->cur[s] + ((size_t)1 << d->b1);
    // part boundaries on multiples of 4096 rows: tiles stay 16-byte aligned
    const size_t per = ((n / d->parts + 4095) & ~size_t(4095));
    HMJ_TRY(dist_wait(ctx, sx, sig_done(s, 0), s, epoch - 1));
    for (int part = 0; part < d->parts; ++part) {
      const size_t lo = std::min(n, per * part);
      const size_t hi = part == d->parts - 1 ? n : std::min(n, per * (part + 1));
      HMJ_TRY(hmj_scatter_level1_dist(ctx, sx, keys + lo, hi - lo, base + (uint32_t)lo, 64 - d->b1, d->b1,
                                      d->cap1, d->peers[s], d->cur[s], flag, flag + 1 + part,
                                      (uint32_t)part * d->ctab, (s == 0 && part == 0) ? 0u : later_ctas));
      HMJ_TRY(dist_signal(ctx, sx, sig_ready(s, part, d->rank)));
    }
  }
  // main stream: the local levels of every part as it arrives, then the join
  hmj_uniform_buckets B[2];
  for (int s = 0; s < 2; ++s)
    for (int part = 0; part < d->parts; ++part) {
      HMJ_TRY(dist_wait(ctx, st, sig_ready(s, part, 0), s, epoch));
      HMJ_TRY(dist_local_levels(ctx, st, s, part, &B[s]));
    }
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_SORTED], st));

  const size_t n_hint = d->rows_total / d->world;
  HMJ_TRY(hmj_bucket_join_run(ctx, HMJ_KEY_U64, B[0], B[1], d->cap_seg,
                              64 - d->b1 - d->b2 - d->b3 - 12, n_hint, n_hint, nullptr, nullptr,
                              nullptr, nullptr, h_num_pairs, ok));
  HMJ_CUDA(ctx, cudaEventRecord(ctx->ev[EV_JOINED], st));
  HMJ_CUDA(ctx, cudaStreamSynchronize(sx));
  if (*ok) ctx->tm.bucket_join = 1;
  return HMJ_OK;
}

// What the last hmj_dist_join_u64 moved, read back from the device: pairs this rank wrote into
// OTHER ranks' memory (level-1 cursors of the regions it does not own, both relations) and the
// pairs of R and of S it received (the owners' count tables).  Synchronises the stream.
extern "C" int hmj_dist_last_exchange(hmj_ctx* ctx, uint64_t* sent_pairs_remote,
                                      uint64_t* received_r, uint64_t* received_s) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->dist != nullptr);
  hmj_dist_state* d = ctx->dist;
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const size_t regions = (size_t)1 << d->b1, subs = (size_t)d->regions_own * d->world;
  std::vector<uint32_t> h(regions > subs ? regions : subs);
  uint64_t sent = 0, recv[2] = {0, 0};
  for (int s = 0; s < 2; ++s) {
    HMJ_CUDA(ctx, cudaMemcpy(h.data(), d->cur[s], regions * 4, cudaMemcpyDeviceToHost));
    for (size_t r = 0; r < regions; ++r)
      if ((int)(r >> d->own_bits) != d->rank) sent += h[r];
    HMJ_CUDA(ctx, cudaMemcpy(h.data(), d->c1[s] + (size_t)(d->parts - 1) * d->ctab, subs * 4, cudaMemcpyDeviceToHost));
    for (size_t r = 0; r < subs; ++r) recv[s] += h[r];
  }
  if (sent_pairs_remote) *sent_pairs_remote = sent;
  if (received_r) *received_r = recv[0];
  if (received_s) *received_s = recv[1];
  return HMJ_OK;
}

extern "C" int hmj_dist_teardown(hmj_ctx* ctx) {
  if (!ctx) return HMJ_EARG;
  hmj_dist_state* d = ctx->dist;
  if (!d) return HMJ_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (d->side_stream) {
    cudaStreamSynchronize(d->side_stream);
    cudaStreamDestroy(d->side_stream);
  }
  if (d->ev_r_sent) cudaEventDestroy(d->ev_r_sent);
  if (d->ev_r_local) cudaEventDestroy(d->ev_r_local);
  for (int r = 0; r < 8; ++r)
    for (int j = 0; j < HANDLES_PER_RANK; ++j)
      if (d->opened[r][j]) cudaIpcCloseMemHandle(d->opened[r][j]);
  for (int s = 0; s < 2; ++s) {
    if (d->k1[s]) cudaFree(d->k1[s]);
    if (d->v1[s]) cudaFree(d->v1[s]);
    if (d->c1[s]) cudaFree(d->c1[s]);
    if (d->cur[s]) cudaFree(d->cur[s]);
  }
  if (d->sig) cudaFree(d->sig);
  delete d;
  ctx->dist = nullptr;
  cudaGetLastError();
  return HMJ_OK;
}
