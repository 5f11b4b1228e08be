// dist_join.cu -- multi-GPU hash-merge join with the exchange FUSED into the first scatter level.
//
// every GPU splits its rows of R and S by hash range, one exchange moves range g
// to GPU g, then every GPU joins its own range.  Here the split and the exchange are one kernel:
// level 1 of the bucket scatter (bucket_scatter.cu) runs over the GPU's local rows, cuts them on
// the top log2(G) hash bits -- exactly partition_only's h >> (64 - bits), partitioned_hash.h:59 --
// and writes every run straight into the owner's memory through NVLink peer pointers (a 4096-pair
// tile leaves as G runs of ~4096/G pairs: KBs per remote store).  Every source GPU owns a
// sub-region of every destination (slot reservation = atomicAdd on a cursor in the SOURCE's
// memory; the counts are handed to the owners after the kernel).  No partition pass, no send
// buffers, no NCCL all-to-all; after a barrier every GPU cuts what it received on the bits below
// in one or two local levels and runs the bucket join.  It stands in for partition_only's
// per-thread push_back into shared partitions (partitioned_hash.h:82-124) at the scale of a node.
// (First plan: level 1 cut on 9 bits, 512 regions spread over the GPUs -- runs of 16 pairs, 128-B
// remote stores, ~300 GB/s of NVLink per GPU; C4 on 8 GPUs 13.4 ms.  GPU bits only: 10.0 ms.)
//
// One process per GPU: the regions are cudaMalloc'ed here and shared with CUDA IPC handles that
// the caller exchanges over torch.distributed (hashmergejoin_b200/dist.py).
#include <math.h>
#include <stdlib.h>

#include "common.cuh"

struct hmj_dist_state {
  int world = 1, rank = 0;
  // level-1 bits (job wide: the GPU bits + own_bits), local level bits (b3 = 0: one local level)
  int g_bits = 0, b1 = 0, b2 = 0, b3 = 0, own_bits = 0;
  uint32_t cap1 = 0, cap2 = 0, cap_seg = 0, regions_own = 0;
  // own level-1 regions per side (0 = R, 1 = S): regions_own x world sub-regions of cap1 slots
  // (keys, row ids) and their counts [regions_own * world], all shared with the peers;
  uint64_t* k1[2] = {nullptr, nullptr};
  uint32_t* v1[2] = {nullptr, nullptr};
  uint32_t* c1[2] = {nullptr, nullptr};
  // private: this source's cursors over all 2^b1 regions of the job (+ 8 words: [2^b1] = flag)
  uint32_t* cur[2] = {nullptr, nullptr};
  hmj_peer_regions peers[2];
  void* opened[2][8][3] = {};
  size_t rows_total = 0;
};

namespace {
constexpr int HANDLE_BYTES = 64;  // sizeof(cudaIpcMemHandle_t)
static_assert(sizeof(cudaIpcMemHandle_t) == HANDLE_BYTES, "IPC handle size");
}  // namespace

// Plans the job for `rows_total` rows per relation over `world` GPUs, allocates this rank's
// level-1 regions and writes their 6 IPC handles (R keys, R row ids, R cursors, S ...) to
// handles_out[6 * 64].
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
  d->b1 = d->g_bits;
  if (const char* e = getenv("HMJ_DIST_L1_BITS")) {
    const int v = atoi(e);
    if (v > d->b1 && v <= 9 && v <= total_bits - 1) d->b1 = v;
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
  for (int s = 0; s < 2; ++s) {
    const size_t subs = (size_t)d->regions_own * world;
    const size_t slots = subs * d->cap1;
    HMJ_CUDA(ctx, cudaMalloc(&d->k1[s], slots * 8));
    HMJ_CUDA(ctx, cudaMalloc(&d->v1[s], slots * 4));
    HMJ_CUDA(ctx, cudaMalloc(&d->c1[s], (subs + 8) * 4));
    HMJ_CUDA(ctx, cudaMemset(d->c1[s], 0, (subs + 8) * 4));
    HMJ_CUDA(ctx, cudaMalloc(&d->cur[s], (((size_t)1 << d->b1) + 8) * 4));
    void* ptrs[3] = {d->k1[s], d->v1[s], d->c1[s]};
    for (int j = 0; j < 3; ++j) {
      cudaIpcMemHandle_t h;
      HMJ_CUDA(ctx, cudaIpcGetMemHandle(&h, ptrs[j]));
      memcpy(handles_out + (size_t)(s * 3 + j) * HANDLE_BYTES, &h, HANDLE_BYTES);
    }
  }
  ctx->dist = d;
  return HMJ_OK;
}

// all_handles[world][6][64]: what every rank wrote in hmj_dist_setup, in rank order.
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
    for (int r = 0; r < d->world; ++r) {
      void* ptrs[3];
      if (r == d->rank) {
        ptrs[0] = d->k1[s]; ptrs[1] = d->v1[s]; ptrs[2] = d->c1[s];
      } else {
        for (int j = 0; j < 3; ++j) {
          cudaIpcMemHandle_t h;
          memcpy(&h, all_handles + ((size_t)r * 6 + s * 3 + j) * HANDLE_BYTES, HANDLE_BYTES);
          HMJ_CUDA(ctx, cudaIpcOpenMemHandle(&ptrs[j], h, cudaIpcMemLazyEnablePeerAccess));
          d->opened[s][r][j] = ptrs[j];
        }
      }
      d->peers[s].keys[r] = reinterpret_cast<uint64_t*>(ptrs[0]);
      d->peers[s].vals[r] = reinterpret_cast<uint32_t*>(ptrs[1]);
      d->peers[s].counts[r] = reinterpret_cast<uint32_t*>(ptrs[2]);
    }
  }
  return HMJ_OK;
}

// Zeroes this rank's cursors and flags (private to the rank: no barrier is needed before the
// scatter; the one after it publishes runs and counts to their owners).
extern "C" int hmj_dist_begin(hmj_ctx* ctx) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->dist != nullptr);
  hmj_dist_state* d = ctx->dist;
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  for (int s = 0; s < 2; ++s)
    HMJ_TRY(hmj_fill_async(ctx, d->cur[s], 0, (((size_t)1 << d->b1) + 8) * 4, ctx->stream));
  return HMJ_OK;
}

// Level 1 over this rank's rows of relation `side` (global row ids rowid_base + i), written into
// the regions of all ranks.  Returns after the kernel has completed (its peer writes are then
// visible to their owners once the caller's barrier has passed).
extern "C" int hmj_dist_scatter(hmj_ctx* ctx, int side, const uint64_t* d_keys, size_t n,
                                uint32_t rowid_base) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->dist != nullptr && (side == 0 || side == 1) && (n == 0 || d_keys));
  hmj_dist_state* d = ctx->dist;
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  uint32_t* flag = d->cur[side] + ((size_t)1 << d->b1);
  HMJ_TRY(hmj_scatter_level1_dist(ctx, d_keys, n, rowid_base, 64 - d->b1, d->b1, d->cap1,
                                  d->peers[side], d->cur[side], flag));
  HMJ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HMJ_OK;
}

// Level 2 + bucket join on this rank's hash range.  *ok = 0: some region or bucket overflowed on
// this rank (the caller agrees on that across ranks and falls back to the NCCL path).
extern "C" int hmj_dist_join(hmj_ctx* ctx, uint64_t* h_num_pairs, int* ok) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->dist != nullptr && h_num_pairs && ok);
  hmj_dist_state* d = ctx->dist;
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  *ok = 0;
  *h_num_pairs = 0;
  const uint32_t regions2 = d->regions_own << d->b2;        // after the first local level
  const uint32_t buckets = regions2 << d->b3;               // final
  const size_t n_hint = d->rows_total / d->world;
  hmj_uniform_buckets B[2];
  for (int s = 0; s < 2; ++s) {
    // scratch words: [8 spare][final cursors: buckets][middle cursors: regions2]
    uint32_t* words = nullptr;
    const size_t nwords = (size_t)8 + buckets + (d->b3 ? regions2 : 0);
    HMJ_TRY(hmj_ws(ctx, s ? WS_US_CURSORS_1 : WS_US_CURSORS, nwords, &words));
    HMJ_TRY(hmj_fill_async(ctx, words, 0, nwords * 4, st));
    uint64_t* k2 = nullptr;
    uint32_t* v2 = nullptr;
    const size_t slots2 = (size_t)buckets * d->cap_seg;
    HMJ_TRY(hmj_ws(ctx, s ? WS_SORT_KEYS_C : WS_SORT_KEYS_B, slots2, &k2));
    HMJ_TRY(hmj_ws(ctx, s ? WS_SORT_VALS_C : WS_SORT_VALS_B, slots2, &v2));
    uint32_t* flag = d->cur[s] + ((size_t)1 << d->b1);  // this rank's level-1 overflows land here too
    const uint32_t subs = d->regions_own * (uint32_t)d->world;
    if (d->b3 == 0) {
      HMJ_TRY(hmj_scatter_level2(ctx, d->k1[s], d->v1[s], d->c1[s], d->cap1, subs, n_hint,
                                 64 - d->b1 - d->b2, d->b2, k2, v2, d->cap_seg, words + 8, flag, d->g_bits));
    } else {
      // sub-regions -> regions2 middle regions (one buffer, used by R then by S) -> final buckets
      uint64_t* km = nullptr;
      uint32_t* vm = nullptr;
      uint32_t* cur_mid = words + 8 + buckets;
      HMJ_TRY(hmj_ws(ctx, WS_SORT_KEYS_A, (size_t)regions2 * d->cap2, &km));
      HMJ_TRY(hmj_ws(ctx, WS_SORT_VALS_A, (size_t)regions2 * d->cap2, &vm));
      HMJ_TRY(hmj_scatter_level2(ctx, d->k1[s], d->v1[s], d->c1[s], d->cap1, subs, n_hint,
                                 64 - d->b1 - d->b2, d->b2, km, vm, d->cap2, cur_mid, flag, d->g_bits));
      HMJ_TRY(hmj_scatter_level2(ctx, km, vm, cur_mid, d->cap2, regions2, n_hint,
                                 64 - d->b1 - d->b2 - d->b3, d->b3, k2, v2, d->cap_seg, words + 8, flag, 0));
    }
    B[s].keys = k2;
    B[s].vals = v2;
    B[s].counts = words + 8;
    B[s].offsets = nullptr;
    B[s].flag = flag;
    B[s].buckets = buckets;
  }
  return hmj_bucket_join_run(ctx, HMJ_KEY_U64, B[0], B[1], d->cap_seg, 64 - d->b1 - d->b2 - d->b3 - 12,
                             n_hint, n_hint, nullptr, nullptr, nullptr, nullptr, h_num_pairs, ok);
}

extern "C" int hmj_dist_teardown(hmj_ctx* ctx) {
  if (!ctx) return HMJ_EARG;
  hmj_dist_state* d = ctx->dist;
  if (!d) return HMJ_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (int s = 0; s < 2; ++s) {
    for (int r = 0; r < 8; ++r)
      for (int j = 0; j < 3; ++j)
        if (d->opened[s][r][j]) cudaIpcCloseMemHandle(d->opened[s][r][j]);
    if (d->k1[s]) cudaFree(d->k1[s]);
    if (d->v1[s]) cudaFree(d->v1[s]);
    if (d->c1[s]) cudaFree(d->c1[s]);
    if (d->cur[s]) cudaFree(d->cur[s]);
  }
  delete d;
  ctx->dist = nullptr;
  return HMJ_OK;
}
