// hash_table.cu -- K7: the build/probe hash-join baseline, kept for comparison with the
// sort-merge path.  Stands in for the loops of BM_hash_join_raw (/root/reference
// hashjoin_bench.cc:47-53):
//     for s in S: s_map[s.key] = s.value          (duplicate keys: the last row wins)
//     for r in R: sum += r.value + s_map[r.key]   (a miss contributes r.value + 0)
// as an open-addressing, linear-probing table in HBM: 16-byte slots {u64 hash, u32 row+1},
// claimed with atomicCAS on the hash word; "last row wins" = atomicMax on the row word.
// Capacity is the power of two >= 2 x |S|.  String keys are verified byte-wise on every
// hash hit.  Algorithmic bytes: build reads 8 (+key) per row and touches one 16 B slot;
// probe reads 8 (+key) per row and >= one 32 B sector.
#include "common.cuh"
#include "join_device.cuh"

namespace {

constexpr uint64_t HT_EMPTY = ~0ull;
constexpr int HT_THREADS = 256;

struct HtSlot {
  unsigned long long hash;
  uint32_t row1;  // row id + 1 (0 = not yet published)
  uint32_t pad;
};

struct HtArgs {
  HtSlot* table;
  uint64_t mask;
  int shift;  // 64 - log2(capacity)
  const uint64_t* hash;
  const uint32_t* rowid;  // null = iota
  uint32_t n;
  const uint64_t* off;
  const uint8_t* chars;
  // the other side's key storage (probe compares against build keys)
  const uint64_t* b_hash_by_row;  // unused
  const uint64_t* b_off;
  const uint8_t* b_chars;
  const uint64_t* vals;    // this side's payloads (null = row id)
  const uint64_t* b_vals;  // build side payloads
  uint64_t* out;           // [0] hits, [1] sum
};

// full-key check of a slot whose hash matched: the word-wise comparison the join kernels use
// (join_device.cuh), 8 key bytes per step whatever the two keys' alignment
__device__ __forceinline__ bool str_rows_equal(const uint64_t* aoff, const uint8_t* ac, uint32_t ar,
                                               const uint64_t* boff, const uint8_t* bc, uint32_t br) {
  return keys_equal_str(aoff, ac, ar, boff, bc, br);
}

__device__ __forceinline__ uint32_t wait_row(const HtSlot* s) {
  uint32_t r;
  do {
    r = *reinterpret_cast<const volatile uint32_t*>(&s->row1);
  } while (r == 0);
  return r - 1;
}

template <int KIND>
__global__ void __launch_bounds__(HT_THREADS) ht_build_kernel(const HtArgs A) {
  const uint32_t stride = gridDim.x * HT_THREADS;
  for (uint32_t i = blockIdx.x * HT_THREADS + threadIdx.x; i < A.n; i += stride) {
    uint64_t h = A.hash[i];
    const uint32_t row = A.rowid ? A.rowid[i] : i;
    if (h == HT_EMPTY) h = HT_EMPTY - 1;  // the sentinel itself cannot be stored
    uint64_t slot = (h * 0x9E3779B97F4A7C15ull) >> A.shift;
    while (true) {
      HtSlot* s = A.table + slot;
      unsigned long long cur = s->hash;
      if (cur == HT_EMPTY) cur = atomicCAS(&s->hash, HT_EMPTY, (unsigned long long)h);
      if (cur == HT_EMPTY) {  // claimed
        atomicMax(&s->row1, row + 1);
        break;
      }
      if (cur == h) {
        bool same = true;
        if (KIND == HMJ_KEY_STRING) {
          const uint32_t other = wait_row(s);
          same = str_rows_equal(A.off, A.chars, row, A.off, A.chars, other);
        } else if (h == HT_EMPTY - 1) {
          const uint32_t other = wait_row(s);
          same = A.hash[other] == A.hash[i];  // rowid==null assumed for this rare path
        }
        if (same) {
          atomicMax(&s->row1, row + 1);  // last row wins
          break;
        }
      }
      slot = (slot + 1) & A.mask;
    }
  }
}

template <int KIND>
__global__ void __launch_bounds__(HT_THREADS) ht_probe_kernel(const HtArgs A) {
  __shared__ unsigned long long s_acc[2];
  if (threadIdx.x < 2) s_acc[threadIdx.x] = 0;
  __syncthreads();
  unsigned long long hits = 0, sum = 0;
  const uint32_t stride = gridDim.x * HT_THREADS;
  for (uint32_t i = blockIdx.x * HT_THREADS + threadIdx.x; i < A.n; i += stride) {
    uint64_t h = A.hash[i];
    const uint32_t row = A.rowid ? A.rowid[i] : i;
    sum += A.vals ? A.vals[row] : (uint64_t)row;
    const uint64_t h_true = h;
    if (h == HT_EMPTY) h = HT_EMPTY - 1;
    uint64_t slot = (h * 0x9E3779B97F4A7C15ull) >> A.shift;
    while (true) {
      const HtSlot* s = A.table + slot;
      const uint64_t cur = s->hash;
      if (cur == HT_EMPTY) break;  // miss: contributes 0
      if (cur == h) {
        const uint32_t brow = s->row1 - 1;
        bool same = true;
        if (KIND == HMJ_KEY_STRING) same = str_rows_equal(A.off, A.chars, row, A.b_off, A.b_chars, brow);
        else if (h == HT_EMPTY - 1) same = A.b_hash_by_row[brow] == h_true;
        if (same) {
          ++hits;
          sum += A.b_vals ? A.b_vals[brow] : (uint64_t)brow;
          break;
        }
      }
      slot = (slot + 1) & A.mask;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    hits += __shfl_down_sync(0xffffffffu, hits, o);
    sum += __shfl_down_sync(0xffffffffu, sum, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&s_acc[0], hits);
    atomicAdd(&s_acc[1], sum);
  }
  __syncthreads();
  if (threadIdx.x < 2 && s_acc[threadIdx.x])
    atomicAdd(reinterpret_cast<unsigned long long*>(A.out) + threadIdx.x, s_acc[threadIdx.x]);
}

}  // namespace

extern "C" int hmj_hashjoin_build_probe(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s,
                                        const uint64_t* d_rvals, const uint64_t* d_svals,
                                        uint64_t* h_hits, uint64_t* h_sum) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, r && s && h_hits && h_sum);
  HMJ_ARG(ctx, r->key_kind == s->key_kind);
  HMJ_ARG(ctx, r->n < (size_t(1) << 32) - 1 && s->n < (size_t(1) << 31));
  HMJ_ARG(ctx, (r->n == 0 || r->d_hash) && (s->n == 0 || s->d_hash));
  // unsorted relations in row order only: a hash equal to the table's empty marker is told apart
  // by reading the hash column at the ROW id, which an explicit row-id column would break
  HMJ_ARG(ctx, r->d_rowid == nullptr && s->d_rowid == nullptr);
  if (r->key_kind == HMJ_KEY_STRING)
    HMJ_ARG(ctx, (r->n == 0 || (r->d_offsets && r->d_chars)) && (s->n == 0 || (s->d_offsets && s->d_chars)));
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  int log2cap = 4;
  while ((size_t(1) << log2cap) < 2 * s->n) ++log2cap;
  const size_t cap = size_t(1) << log2cap;
  HtSlot* table = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_TABLE, cap, &table));
  // EMPTY hash = all ones, row1 must be 0: two memsets over the interleaved slots would be
  // wasteful, so fill with 0xFF and let row1 use the complement convention instead.
  HMJ_TRY(hmj_fill_async(ctx, table, 0xFF, cap * sizeof(HtSlot), st));
  // row1 == 0xFFFFFFFF after the fill; a tiny kernel-free trick is not possible, so reset row1
  // with a strided 2D memset (4 bytes every 16).
  HMJ_CUDA(ctx, cudaMemset2DAsync(reinterpret_cast<char*>(table) + 8, sizeof(HtSlot), 0, 4, cap, st));
  uint64_t* scal = nullptr;
  HMJ_TRY(hmj_ws(ctx, WS_SCALARS, 8, &scal));
  HMJ_TRY(hmj_fill_async(ctx, scal, 0, 16, st));

  HtArgs B = {};
  B.table = table; B.mask = cap - 1; B.shift = 64 - log2cap;
  B.hash = s->d_hash; B.rowid = s->d_rowid; B.n = (uint32_t)s->n;
  B.off = s->d_offsets; B.chars = s->d_chars;
  HtArgs P = {};
  P.table = table; P.mask = cap - 1; P.shift = 64 - log2cap;
  P.hash = r->d_hash; P.rowid = r->d_rowid; P.n = (uint32_t)r->n;
  P.off = r->d_offsets; P.chars = r->d_chars;
  P.b_off = s->d_offsets; P.b_chars = s->d_chars; P.b_hash_by_row = s->d_hash;
  P.vals = d_rvals; P.b_vals = d_svals; P.out = scal;
  const int gb = (int)min(hmj_div_up(s->n, HT_THREADS * 4), (size_t)ctx->num_sms * 8);
  const int gp = (int)min(hmj_div_up(r->n, HT_THREADS * 4), (size_t)ctx->num_sms * 8);
  if (s->n) {
    if (r->key_kind == HMJ_KEY_U64) ht_build_kernel<HMJ_KEY_U64><<<gb, HT_THREADS, 0, st>>>(B);
    else ht_build_kernel<HMJ_KEY_STRING><<<gb, HT_THREADS, 0, st>>>(B);
    HMJ_LAUNCHED(ctx);
  }
  if (r->n) {
    if (r->key_kind == HMJ_KEY_U64) ht_probe_kernel<HMJ_KEY_U64><<<gp, HT_THREADS, 0, st>>>(P);
    else ht_probe_kernel<HMJ_KEY_STRING><<<gp, HT_THREADS, 0, st>>>(P);
    HMJ_LAUNCHED(ctx);
  }
  uint64_t* h = reinterpret_cast<uint64_t*>(ctx->pinned);
  HMJ_CUDA(ctx, cudaMemcpyAsync(h, scal, 16, cudaMemcpyDeviceToHost, st));
  HMJ_CUDA(ctx, cudaStreamSynchronize(st));
  *h_hits = h[0];
  *h_sum = h[1];
  return HMJ_OK;
}
