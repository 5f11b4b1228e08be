// dist_strings.cu -- the hash-range exchange of STRING keys, fused into the key gather.
//
// "string keys travel with their pairs".  The NCCL route (dist.py) puts a rank's
// keys into send order with a gather kernel, then moves row ids, lengths and key bytes with three
// all_to_all_single calls: on 8 B200 the gather took 6.9 ms and the collectives 7-9 ms per
// relation of 64M rows (profiles/r2n_c5_n8_exchange_phases.txt), ~210 GB/s per GPU.  Here the
// gather kernel writes every key STRAIGHT into its owner's receive buffer over NVLink peer
// pointers (the send-order copy is never materialised locally) and a second small kernel does the
// same for the row ids and lengths: it replaces partition_only's row movement
// (/root/reference partitioned_hash.h:102-110) at node scale.  Where a rank's block lands inside
// an owner's buffer follows from the G x G count matrix every rank already holds.
// One process per GPU; receive buffers are cudaMalloc'ed here once (capacity fixed at setup) and
// shared through CUDA IPC handles exchanged by the caller.  Ordering between the ranks is the
// caller's (a stream-ordered NCCL barrier after the sends and before the buffers are reused).
#include "common.cuh"

namespace {
constexpr int XS_HANDLE_BYTES = 64;
constexpr int XS_HANDLES = 6;  // per side: chars, row ids, lengths

struct xs_peers {
  uint8_t* chars[8];
  uint32_t* rowid[8];
  uint32_t* lens[8];
  // where this rank's block for owner g starts in g's buffers, and the send-order range of owner g
  uint64_t byte_base[8];
  uint64_t row_base[8];
  uint64_t part_first[9];  // rows [part_first[g], part_first[g+1]) of the send order go to owner g
  uint32_t tile_first[9];  // CTAs [tile_first[g], tile_first[g+1]) serve owner g
};

constexpr int XS_THREADS = 256;
constexpr int XS_ROWS = 1024;           // send-order rows per CTA (all for one owner)
constexpr int XS_CHUNK = 32 * 1024;     // key bytes a CTA stages (mean 18 B per key: 18 KB)
constexpr size_t XS_SMEM = 2 * (size_t)(XS_ROWS + 2) * 8 + XS_CHUNK + 32;

// One CTA = up to 1024 consecutive send-order rows of ONE owner.  Their key bytes are contiguous in
// the owner's buffer, so the CTA gathers them (random rows of the local column) into shared
// memory in send order and then writes the whole chunk to the peer with 16-byte coalesced stores;
// row ids and lengths leave as coalesced 4-byte stores.  (A first version let every row's 8
// threads store its bytes straight into peer memory: byte-granular remote stores, 9 ms per 64M
// rows on 2 GPUs against 4.2 + 5.7 ms for gather + NCCL.)
__global__ void __launch_bounds__(XS_THREADS)
xchg_send_kernel(const uint8_t* __restrict__ chars, const uint64_t* __restrict__ start,
                 const uint64_t* __restrict__ send_off, const uint32_t* __restrict__ rowid,
                 const uint32_t* __restrict__ lens, int world, const xs_peers P) {
  extern __shared__ __align__(16) unsigned char xs_raw[];
  uint64_t* s_off = reinterpret_cast<uint64_t*>(xs_raw);                 // [rows + 1]
  uint64_t* s_start = s_off + XS_ROWS + 2;                               // [rows] first byte of a key
  unsigned char* s_buf = xs_raw + 2 * (size_t)(XS_ROWS + 2) * 8;         // [XS_CHUNK + 32]
  const int tid = threadIdx.x;
  int g = 0;
#pragma unroll
  for (int k = 1; k < 8; ++k)
    if (k < world && blockIdx.x >= P.tile_first[k]) g = k;
  const uint64_t i0 = P.part_first[g] + (uint64_t)(blockIdx.x - P.tile_first[g]) * XS_ROWS;
  const uint64_t i1 = min(i0 + (uint64_t)XS_ROWS, P.part_first[g + 1]);
  if (i0 >= i1) return;
  const uint32_t rows = (uint32_t)(i1 - i0);
  for (uint32_t r = tid; r <= rows; r += XS_THREADS) s_off[r] = send_off[i0 + r];
  for (uint32_t r = tid; r < rows; r += XS_THREADS) s_start[r] = start[i0 + r];
  const uint64_t part_off = send_off[P.part_first[g]];
  __syncthreads();
  const uint64_t base = s_off[0], total = s_off[rows] - base;
  uint8_t* dst = P.chars[g] + P.byte_base[g] + (base - part_off);
  // row ids and lengths: consecutive rows, consecutive slots
  {
    const uint64_t r0 = P.row_base[g] + (i0 - P.part_first[g]);
    for (uint32_t r = tid; r < rows; r += XS_THREADS) {
      P.rowid[g][r0 + r] = rowid[i0 + r];
      P.lens[g][r0 + r] = lens[i0 + r];
    }
  }
  if (total <= (uint64_t)XS_CHUNK) {
    // the chunk sits in shared memory at the same offset modulo 16 as in the owner's buffer
    const uint32_t pad = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15);
    // Four rows per 8-lane group at a time, all their byte loads issued before the first store: a
    // row's key sits at a random place of the column, and with one row in flight per group the
    // CTA spent its time waiting for one DRAM round trip after the other.
    const int lane = tid & 7;
    constexpr int GROUPS = XS_THREADS / 8, Q = 4, T = 4;  // T x 8 = 32 bytes per row in registers
    for (uint32_t rb = tid >> 3; rb < rows; rb += GROUPS * Q) {
      uint8_t v[Q][T];
      uint32_t o[Q], len[Q];
      const uint8_t* src[Q];
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const uint32_t r = rb + q * GROUPS;
        len[q] = 0;
        if (r < rows) {
          o[q] = (uint32_t)(s_off[r] - base);
          len[q] = (uint32_t)(s_off[r + 1] - s_off[r]);
          src[q] = chars + s_start[r];
#pragma unroll
          for (int t = 0; t < T; ++t) {
            const uint32_t k = lane + 8 * t;
            v[q][t] = k < len[q] ? src[q][k] : (uint8_t)0;
          }
        }
      }
#pragma unroll
      for (int q = 0; q < Q; ++q) {
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const uint32_t k = lane + 8 * t;
          if (k < len[q]) s_buf[pad + o[q] + k] = v[q][t];
        }
        for (uint32_t k = lane + 8 * T; k < len[q]; k += 8) s_buf[pad + o[q] + k] = src[q][k];
      }
    }
    __syncthreads();
    const uint32_t n = (uint32_t)total;
    const uint32_t head = min(n, (16u - pad) & 15u);
    for (uint32_t k = tid; k < head; k += XS_THREADS) dst[k] = s_buf[pad + k];
    const uint32_t vecs = (n - head) >> 4;
    const uint4* sv = reinterpret_cast<const uint4*>(s_buf + pad + head);
    uint4* dv = reinterpret_cast<uint4*>(dst + head);
    for (uint32_t v = tid; v < vecs; v += XS_THREADS) dv[v] = sv[v];
    for (uint32_t k = head + (vecs << 4) + tid; k < n; k += XS_THREADS) dst[k] = s_buf[pad + k];
  } else {
    // long keys: every row's 8 threads copy it directly
    const int lane = tid & 7;
    for (uint32_t r = tid >> 3; r < rows; r += XS_THREADS / 8) {
      const uint64_t o = s_off[r] - base, len = s_off[r + 1] - s_off[r];
      const uint8_t* src = chars + s_start[r];
      for (uint64_t k = lane; k < len; k += 8) dst[o + k] = src[k];
    }
  }
}
}  // namespace

struct hmj_xchg_state {
  int world = 1, rank = 0;
  size_t rows_cap = 0, bytes_cap = 0;
  uint8_t* chars[2] = {nullptr, nullptr};
  uint32_t* rowid[2] = {nullptr, nullptr};
  uint32_t* lens[2] = {nullptr, nullptr};
  void* peer[2][8][3] = {};
  void* opened[8][XS_HANDLES] = {};
  bool connected = false;
};

extern "C" int hmj_xchg_handle_bytes(void) { return XS_HANDLES * XS_HANDLE_BYTES; }

// Allocates this rank's receive buffers (both relations: rows_cap rows, bytes_cap key bytes each)
// and writes their 6 IPC handles to handles_out[hmj_xchg_handle_bytes()].
extern "C" int hmj_xchg_setup(hmj_ctx* ctx, int rank, int world, size_t rows_cap, size_t bytes_cap,
                              unsigned char* handles_out) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, world >= 2 && world <= 8 && rank >= 0 && rank < world && handles_out && ctx->xchg == nullptr);
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  hmj_xchg_state* x = new hmj_xchg_state();
  x->world = world;
  x->rank = rank;
  x->rows_cap = rows_cap;
  x->bytes_cap = bytes_cap;
  ctx->xchg = x;  // hmj_xchg_teardown frees what has been allocated
  for (int s = 0; s < 2; ++s) {
    HMJ_CUDA(ctx, cudaMalloc(&x->chars[s], bytes_cap + 64));
    HMJ_CUDA(ctx, cudaMalloc(&x->rowid[s], (rows_cap + 16) * 4));
    HMJ_CUDA(ctx, cudaMalloc(&x->lens[s], (rows_cap + 16) * 4));
    void* ptrs[3] = {x->chars[s], x->rowid[s], x->lens[s]};
    for (int j = 0; j < 3; ++j) {
      cudaIpcMemHandle_t h;
      HMJ_CUDA(ctx, cudaIpcGetMemHandle(&h, ptrs[j]));
      memcpy(handles_out + (size_t)(s * 3 + j) * XS_HANDLE_BYTES, &h, XS_HANDLE_BYTES);
    }
  }
  return HMJ_OK;
}

extern "C" int hmj_xchg_connect(hmj_ctx* ctx, const unsigned char* all_handles) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->xchg != nullptr && all_handles != nullptr);
  hmj_xchg_state* x = ctx->xchg;
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  for (int r = 0; r < x->world; ++r)
    for (int s = 0; s < 2; ++s)
      for (int j = 0; j < 3; ++j) {
        void* p = nullptr;
        if (r == x->rank) {
          p = j == 0 ? (void*)x->chars[s] : j == 1 ? (void*)x->rowid[s] : (void*)x->lens[s];
        } else {
          cudaIpcMemHandle_t h;
          memcpy(&h, all_handles + ((size_t)r * XS_HANDLES + s * 3 + j) * XS_HANDLE_BYTES, XS_HANDLE_BYTES);
          HMJ_CUDA(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
          x->opened[r][s * 3 + j] = p;
        }
        x->peer[s][r][j] = p;
      }
  x->connected = true;
  return HMJ_OK;
}

// Sends this rank's n rows of relation `side`, given in send order (grouped by owner):
// d_start[i] = first byte of row i's key in d_chars, d_send_off[n+1] = exclusive prefix sum of the
// lengths, d_rowid / d_lens = what the owner receives per row.  h_part_first[world+1] = send-order
// row range per owner; h_row_base / h_byte_base[world] = where this rank's block starts in the
// owner's buffers.  Enqueues one kernel; nothing is synchronised.
extern "C" int hmj_xchg_send(hmj_ctx* ctx, int side, const uint8_t* d_chars, const uint64_t* d_start,
                             const uint64_t* d_send_off, const uint32_t* d_rowid,
                             const uint32_t* d_lens, size_t n, const uint64_t* h_part_first,
                             const uint64_t* h_row_base, const uint64_t* h_byte_base) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->xchg != nullptr && ctx->xchg->connected && (side == 0 || side == 1));
  if (n == 0) return HMJ_OK;
  HMJ_ARG(ctx, d_start && d_send_off && d_rowid && d_lens && h_part_first && h_row_base && h_byte_base);
  hmj_xchg_state* x = ctx->xchg;
  HMJ_CUDA(ctx, cudaSetDevice(ctx->device));
  xs_peers P = {};
  for (int g = 0; g < x->world; ++g) {
    P.chars[g] = reinterpret_cast<uint8_t*>(x->peer[side][g][0]);
    P.rowid[g] = reinterpret_cast<uint32_t*>(x->peer[side][g][1]);
    P.lens[g] = reinterpret_cast<uint32_t*>(x->peer[side][g][2]);
    P.byte_base[g] = h_byte_base[g];
    P.row_base[g] = h_row_base[g];
    P.part_first[g] = h_part_first[g];
  }
  for (int g = x->world; g <= 8; ++g) P.part_first[g] = h_part_first[x->world];
  uint32_t tiles = 0;
  for (int g = 0; g < 8; ++g) {
    P.tile_first[g] = tiles;
    if (g < x->world) tiles += (uint32_t)hmj_div_up((size_t)(h_part_first[g + 1] - h_part_first[g]), XS_ROWS);
  }
  P.tile_first[8] = tiles;
  cudaEvent_t ev;
  HMJ_TRY(hmj_prof_begin(ctx, HMJ_K_XSCATTER, 0.0, &ev));
  HMJ_CUDA(ctx, cudaFuncSetAttribute(xchg_send_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)XS_SMEM));
  xchg_send_kernel<<<tiles, XS_THREADS, XS_SMEM, ctx->stream>>>(d_chars, d_start, d_send_off, d_rowid, d_lens,
                                                                x->world, P);
  HMJ_LAUNCHED(ctx);
  hmj_prof_end(ctx, ev);
  return HMJ_OK;
}

// this rank's receive buffers of relation `side` and their capacities
extern "C" int hmj_xchg_buffers(hmj_ctx* ctx, int side, uint8_t** d_chars, uint32_t** d_rowid,
                                uint32_t** d_lens, size_t* rows_cap, size_t* bytes_cap) {
  if (!ctx) return HMJ_EARG;
  HMJ_ARG(ctx, ctx->xchg != nullptr && (side == 0 || side == 1));
  hmj_xchg_state* x = ctx->xchg;
  if (d_chars) *d_chars = x->chars[side];
  if (d_rowid) *d_rowid = x->rowid[side];
  if (d_lens) *d_lens = x->lens[side];
  if (rows_cap) *rows_cap = x->rows_cap;
  if (bytes_cap) *bytes_cap = x->bytes_cap;
  return HMJ_OK;
}

extern "C" int hmj_xchg_teardown(hmj_ctx* ctx) {
  if (!ctx) return HMJ_EARG;
  hmj_xchg_state* x = ctx->xchg;
  if (!x) return HMJ_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (int r = 0; r < 8; ++r)
    for (int j = 0; j < XS_HANDLES; ++j)
      if (x->opened[r][j]) cudaIpcCloseMemHandle(x->opened[r][j]);
  for (int s = 0; s < 2; ++s) {
    if (x->chars[s]) cudaFree(x->chars[s]);
    if (x->rowid[s]) cudaFree(x->rowid[s]);
    if (x->lens[s]) cudaFree(x->lens[s]);
  }
  delete x;
  ctx->xchg = nullptr;
  cudaGetLastError();
  return HMJ_OK;
}
