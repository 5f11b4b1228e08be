/*
 * hmj.h -- C ABI of libhmj.so, the B200 (sm_100a) hash -> radix-sort -> merge-join path.
 *
 * This is the drop-in boundary for the hot path of dryman/HashMergeJoin.  The reference has
 * no FFI: its interface is C++11 header templates.  Each entry point below names the reference
 * template/loop it stands in for; the C++ facade headers in hashmergejoin_b200/facade/
 * (radix_hash.h, hashjoin.h, partitioned_hash.h, radix_sort.h) keep the reference's names and
 * argument order and call only these functions.
 *
 * Conventions
 *  - every function returns an int status: HMJ_OK (0) or a negative HMJ_E* code;
 *    hmj_last_error(ctx) gives the text.  There is NO CPU fallback: without a usable CUDA
 *    device hmj_create fails with HMJ_ENODEVICE.
 *  - pointers named d_* are DEVICE pointers (HBM of the context's device); pointers named
 *    h_* are HOST pointers (pageable or pinned).  Sizes are element counts unless "bytes".
 *  - relations are columnar: integer keys as u64[n]; string keys as offsets u64[n+1] into a
 *    chars byte buffer (row i = chars[offsets[i] .. offsets[i+1])).  Row ids are u32, so a
 *    relation has at most 2^32-1 rows (the reference caps n at 2^31-1, radix_hash.h:361).
 *  - all calls are asynchronous on the context's stream unless they return a scalar to the
 *    host (then they synchronise that stream).  One context = one device + one stream;
 *    a context is not thread-safe.
 */
#ifndef HMJ_H
#define HMJ_H 1

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HMJ_OK 0
#define HMJ_EARG (-1)       /* bad argument */
#define HMJ_ECUDA (-2)      /* CUDA runtime error */
#define HMJ_ENOMEM (-3)     /* device or host allocation failed */
#define HMJ_ENODEVICE (-4)  /* no CUDA device / wrong architecture */
#define HMJ_ECAPACITY (-5)  /* caller-provided output capacity too small */

typedef struct hmj_ctx hmj_ctx;

/* key description used by the join / hash-table entry points */
#define HMJ_KEY_U64 0     /* std::hash<integral> = identity: hash equality == key equality */
#define HMJ_KEY_STRING 1  /* std::hash<std::string>: verify key bytes on every hash match */

/* ---- context ------------------------------------------------------------------------- */
/* Creates a context on `device` with its own non-blocking stream. */
int hmj_create(hmj_ctx** ctx, int device);
/* Same, but work is issued on the caller's cudaStream_t (e.g. torch's current stream). */
int hmj_create_on_stream(hmj_ctx** ctx, int device, void* cuda_stream);
int hmj_destroy(hmj_ctx* ctx);
const char* hmj_last_error(hmj_ctx* ctx);
int hmj_synchronize(hmj_ctx* ctx);
/* library/version probe usable without a GPU (returns e.g. "hmj-b200 0.1 sm_100a"). */
const char* hmj_version(void);
/* number of kernels launched by this context since creation (for bench gpu_launches). */
int hmj_launch_count(hmj_ctx* ctx, uint64_t* count);
/* releases the cached workspace (scratch buffers grow on demand and are reused). */
int hmj_release_workspace(hmj_ctx* ctx);

/* ---- A9: radix width helper ---------------------------------------------------------- */
/* replaces radix_hash::optimal_partition (radix_hash.h:38-57); host scalar, no GPU needed. */
int hmj_optimal_partition(size_t input_num);

/* ---- A1: hashing --------------------------------------------------------------------- */
/* replaces Hash{}(key) with Hash = std::hash<integral> (identity) at radix_hash.h:314,340 */
int hmj_hash_u64(hmj_ctx* ctx, const uint64_t* d_keys, size_t n, uint64_t* d_hash);
/* replaces Hash{}(key) with Hash = std::hash<std::string> (libstdc++ _Hash_bytes, seed
 * 0xc70f6907) at radix_hash.h:314,340 and partitioned_hash.h:58,76 */
int hmj_hash_strings(hmj_ctx* ctx, const uint64_t* d_offsets, const uint8_t* d_chars,
                     size_t n, uint64_t* d_hash);

/* ---- A2+A3+A4: sort of (hash,row-id) pairs ------------------------------------------- */
/* replaces radix_hash::radix_inplace_seq / radix_inplace_par (radix_hash.h:425-654) on a
 * pre-hashed column.  Sorts d_hash ascending, permuting d_rowid alongside; the sort is stable,
 * so equal hashes keep ascending input order.  d_rowid == NULL on input means "iota": pass
 * d_rowid_out to receive the permutation.  Results are left in d_hash_out / d_rowid_out
 * (may alias the inputs). */
int hmj_sort_pairs(hmj_ctx* ctx, const uint64_t* d_hash, const uint32_t* d_rowid, size_t n,
                   uint64_t* d_hash_out, uint32_t* d_rowid_out);
/* replaces radix_hash::radix_non_inplace_par<Key,Value,std::hash<Key>> (radix_hash.h:351-423):
 * hash + sort in one call; row id = input index. */
int hmj_hash_sort_u64(hmj_ctx* ctx, const uint64_t* d_keys, size_t n, uint64_t* d_hash_out,
                      uint32_t* d_rowid_out);
int hmj_hash_sort_strings(hmj_ctx* ctx, const uint64_t* d_offsets, const uint8_t* d_chars,
                          size_t n, uint64_t* d_hash_out, uint32_t* d_rowid_out);
/* replaces ::radix_int_inplace / ::radix_int_non_inplace (radix_sort.h:333-398,452-522):
 * sort of (unsigned key, u32 payload) pairs without hashing.  Same kernel as hmj_sort_pairs. */
int hmj_sort_int_pairs(hmj_ctx* ctx, const uint64_t* d_keys, const uint32_t* d_vals, size_t n,
                       uint64_t* d_keys_out, uint32_t* d_vals_out);

/* ---- A5: merge join of two sorted (hash,row-id) streams ------------------------------ */
/* A relation as the join sees it: the sorted pairs plus (for strings) the key storage the
 * row ids point into, used for full-key verification on every hash match. */
typedef struct hmj_relation {
  const uint64_t* d_hash;    /* sorted ascending, n entries */
  const uint32_t* d_rowid;   /* row id of each sorted entry */
  size_t n;
  int key_kind;              /* HMJ_KEY_U64 or HMJ_KEY_STRING */
  const uint64_t* d_offsets; /* HMJ_KEY_STRING: n+1 offsets (indexed by row id) */
  const uint8_t* d_chars;    /* HMJ_KEY_STRING: key bytes */
} hmj_relation;

/* replaces the merge walk of HashMergeJoin::iterator (hashjoin.h:70-180) as a TRUE equi-join:
 * every (r,s) with equal hash and equal key.  Pass 1: exact number of matching pairs. */
int hmj_merge_join_count(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s,
                         uint64_t* h_num_pairs);
/* Pass 2: writes the pairs as (build_row = R row id, probe_row = S row id) into two u32
 * device arrays of capacity `cap` pairs; *h_num_pairs receives the count written.
 * Returns HMJ_ECAPACITY (nothing written beyond cap) when cap is too small. */
int hmj_merge_join_write(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s,
                         uint32_t* d_out_r, uint32_t* d_out_s, size_t cap,
                         uint64_t* h_num_pairs);
/* the consumer of hashjoin_bench.cc:131-133: count and sum += r.value + s.value over all
 * matches, without materialising pairs.  d_rvals/d_svals are u64 payload columns indexed by
 * row id; NULL means "payload = row id". */
int hmj_merge_join_aggregate(hmj_ctx* ctx, const hmj_relation* r, const hmj_relation* s,
                             const uint64_t* d_rvals, const uint64_t* d_svals,
                             uint64_t* h_num_pairs, uint64_t* h_sum);

/* ---- whole operator: HashMergeJoin(r_begin,r_end,s_begin,s_end) ----------------------- */
/* replaces HashMergeJoin ctor + full iteration (hashjoin.h:56-68,183-195) with DEVICE inputs.
 * Same result multiset as hash -> sort -> merge walk.  Two plans, chosen by the library:
 *  (a) radix-partition + in-shared-memory bucket join: both sides are cut on the same top hash
 *      bits (bucket_scatter.cu), every bucket pair is joined in shared memory (bucket_join.cu);
 *      nothing is sorted or merged.  Runs when the hashes spread evenly (hmj_timings.bucket_join
 *      = 1).  Pair order: bucket arrival order (unspecified; HMJ_JOIN_ORDER=hash gives ascending
 *      bucket order, deterministic between runs).
 *  (b) onesweep sort of both sides + merge-path join (radix_sort.cu, merge_join.cu) for skewed
 *      or heavily duplicated hashes; pairs come in ascending hash order.
 * The result stays in context-owned device buffers until the next join on this context.
 * *h_num_pairs receives the pair count. */
int hmj_join_u64(hmj_ctx* ctx, const uint64_t* d_rkeys, size_t nr, const uint64_t* d_skeys,
                 size_t ns, uint64_t* h_num_pairs);
/* The local step of the multi-GPU join: two UNSORTED relations of (hash, global
 * row id) pairs as they arrive from the hash-range exchange.  All hashes share their topmost
 * `prefix_bits` bits (the rank's hash range, partitioned_hash.h:59 h >> (64 - bits)); identity
 * hash, so equal hashes are equal keys.  Result as hmj_join_u64 (hmj_join_result*). */
int hmj_join_pairs(hmj_ctx* ctx, const uint64_t* d_rhash, const uint32_t* d_rrowid, size_t nr,
                   const uint64_t* d_shash, const uint32_t* d_srowid, size_t ns, int prefix_bits,
                   uint64_t* h_num_pairs);
int hmj_join_strings(hmj_ctx* ctx, const uint64_t* d_roffsets, const uint8_t* d_rchars,
                     size_t nr, const uint64_t* d_soffsets, const uint8_t* d_schars, size_t ns,
                     uint64_t* h_num_pairs);
/* hmj_join_strings on one hash range: every key's std::hash shares its topmost `prefix_bits` bits
 * (the rank's range after the hash-range exchange of the multi-GPU join, partitioned_hash.h:59
 * h >> (64 - bits)).  The bucket plan skips those bits; results are the same as with 0. */
int hmj_join_strings_range(hmj_ctx* ctx, const uint64_t* d_roffsets, const uint8_t* d_rchars,
                           size_t nr, const uint64_t* d_soffsets, const uint8_t* d_schars,
                           size_t ns, int prefix_bits, uint64_t* h_num_pairs);
/* Copies the keys of rows d_rows[0..n) into a new column pair in that order: d_out_offsets[n+1]
 * is given (exclusive prefix sum of the rows' lengths), d_out_chars receives the bytes.  The
 * string side of partition_only's row movement (partitioned_hash.h:102-110): keys are put in
 * send order once, before the hash-range exchange. */
int hmj_gather_strings(hmj_ctx* ctx, const uint64_t* d_offsets, const uint8_t* d_chars,
                       const uint32_t* d_rows, size_t n, const uint64_t* d_out_offsets,
                       uint8_t* d_out_chars);
/* Order of the result pairs of the whole-operator joins on this context.  0 (default): whatever
 * the faster plan yields (bucket arrival order on the bucket path, not reproducible between
 * runs).  1: ascending hash order, equal hashes by (build row, probe row) position in the sorted
 * relations -- the order HashMergeJoin::iterator walks in (hashjoin.h:126-153), identical between
 * runs; the join then always takes the sort + merge-path plan. */
int hmj_set_join_order(hmj_ctx* ctx, int hash_order);
/* The same in two steps for callers that also need the rows' lengths (the hash-range exchange
 * sends them): row_spans reads every row's offset pair once -> start byte and length per row;
 * after an exclusive prefix sum of the lengths (d_out_offsets), gather_spans copies the bytes. */
int hmj_string_row_spans(hmj_ctx* ctx, const uint64_t* d_offsets, const uint32_t* d_rows, size_t n,
                         uint64_t* d_start_out, uint32_t* d_len_out);
int hmj_gather_spans(hmj_ctx* ctx, const uint8_t* d_chars, const uint64_t* d_start, size_t n,
                     const uint64_t* d_out_offsets, uint8_t* d_out_chars);
/* device pointers to the last join's pairs (valid until the next join / destroy). */
int hmj_join_result(hmj_ctx* ctx, const uint32_t** d_out_r, const uint32_t** d_out_s,
                    uint64_t* num_pairs);
/* copies the last join's pairs to host arrays of capacity cap. */
int hmj_join_result_to_host(hmj_ctx* ctx, uint32_t* h_out_r, uint32_t* h_out_s, size_t cap);
/* Same, into caller-owned device buffers (stream-ordered, not synchronised): for callers that
 * keep working on the device, e.g. the multi-GPU join mapping arrival positions to global rows. */
int hmj_join_result_copy(hmj_ctx* ctx, uint32_t* d_out_r, uint32_t* d_out_s, size_t cap);

/* Same operator with HOST buffers: H2D copies, join and D2H of the pairs inside the call.
 * This is what the C++ facade and the e2e benchmark call. h_out_* may be NULL (count only). */
int hmj_join_u64_host(hmj_ctx* ctx, const uint64_t* h_rkeys, size_t nr, const uint64_t* h_skeys,
                      size_t ns, uint32_t* h_out_r, uint32_t* h_out_s, size_t cap,
                      uint64_t* h_num_pairs);
int hmj_join_strings_host(hmj_ctx* ctx, const uint64_t* h_roffsets, const uint8_t* h_rchars,
                          size_t nr, const uint64_t* h_soffsets, const uint8_t* h_schars,
                          size_t ns, uint32_t* h_out_r, uint32_t* h_out_s, size_t cap,
                          uint64_t* h_num_pairs);
/* host-buffer versions of the sort entry points (facade: radix_non_inplace_par etc.). */
int hmj_hash_sort_u64_host(hmj_ctx* ctx, const uint64_t* h_keys, size_t n, uint64_t* h_hash_out,
                           uint32_t* h_rowid_out);
int hmj_hash_sort_strings_host(hmj_ctx* ctx, const uint64_t* h_offsets, const uint8_t* h_chars,
                               size_t n, uint64_t* h_hash_out, uint32_t* h_rowid_out);
int hmj_sort_pairs_host(hmj_ctx* ctx, const uint64_t* h_hash, const uint32_t* h_rowid, size_t n,
                        uint64_t* h_hash_out, uint32_t* h_rowid_out);

/* Page-locked host memory for callers that stage columns themselves (the C++ facade flattens the
 * caller's containers into such arrays): copies between them and the GPU run at the host link's
 * rate and asynchronously.  No context needed. */
int hmj_host_alloc(void** out, size_t bytes);
int hmj_host_free(void* p);

/* ---- A6: partitioning on the top hash bits ------------------------------------------- */
/* replaces the counting loop of partition_only (partitioned_hash.h:57-60): counts[p] = number
 * of rows with (hash >> (64-bits)) == p; d_counts has 1<<bits u64 entries. bits in [1,16]. */
int hmj_partition_counts(hmj_ctx* ctx, const uint64_t* d_hash, size_t n, int bits,
                         uint64_t* d_counts);
/* replaces the scatter of partition_only (partitioned_hash.h:75-79), deterministically:
 * rows are grouped by partition, stable inside a partition.  d_rowid NULL = iota.
 * d_part_offsets receives (1<<bits)+1 u64 exclusive offsets.  bits in [1,8]. */
int hmj_partition_pairs(hmj_ctx* ctx, const uint64_t* d_hash, const uint32_t* d_rowid, size_t n,
                        int bits, uint64_t* d_hash_out, uint32_t* d_rowid_out,
                        uint64_t* d_part_offsets);
/* same, with row ids generated as rowid_base + input index when d_rowid is NULL: the multi-GPU
 * path gives every rank's rows globally unique ids this way. */
int hmj_partition_pairs_base(hmj_ctx* ctx, const uint64_t* d_hash, const uint32_t* d_rowid,
                             uint32_t rowid_base, size_t n, int bits, uint64_t* d_hash_out,
                             uint32_t* d_rowid_out, uint64_t* d_part_offsets);
int hmj_partition_counts_host(hmj_ctx* ctx, const uint64_t* h_hash, size_t n, int bits,
                              uint64_t* h_counts);
/* hmj_partition_pairs_base for callers that may not allocate device memory (code under CUDA
 * graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
 * reference.  Same arguments, same result; the
 * scratch (digit histogram, tile counter, look-back descriptors) lives in a fixed 8.01 MiB
 * array of device memory inside the library (held by every GPU that loads the library until the
 * process exits), so n is limited to HMJ_NOALLOC_MAX_ROWS: a larger n fails with HMJ_EARG and a
 * message naming the limit.  The call launches kernels on the context's stream and calls no
 * cudaMalloc / cudaFree and no allocating library routine; the first call per GPU looks up the
 * array's address (make it -- n = 0 is enough -- before a region that must not reach the
 * driver).  All calls on one GPU share the array: each clears what it needs at zero, and calls
 * that overlap on the GPU (two contexts or streams) race -- serialise them. */
#define HMJ_NOALLOC_MAX_ROWS ((size_t)1 << 24)
int hmj_partition_pairs_noalloc(hmj_ctx* ctx, const uint64_t* d_hash, const uint32_t* d_rowid,
                                uint32_t rowid_base, size_t n, int bits, uint64_t* d_hash_out,
                                uint32_t* d_rowid_out, uint64_t* d_part_offsets);

/* ---- A7: hash-join baseline (build/probe) -------------------------------------------- */
/* replaces the loops of BM_hash_join_raw (hashjoin_bench.cc:47-53): build an open-addressing
 * table over S (duplicate keys: last row wins, as s_map[k]=v), probe with every R row and
 * return sum += r.value + (hit ? s.value : 0) and the number of hits.  d_*vals NULL = row id.
 * Both relations in row order (d_rowid must be NULL). */
int hmj_hashjoin_build_probe(hmj_ctx* ctx, const hmj_relation* r_unsorted,
                             const hmj_relation* s_unsorted, const uint64_t* d_rvals,
                             const uint64_t* d_svals, uint64_t* h_hits, uint64_t* h_sum);

/* ---- multi-GPU join with the exchange fused into the first scatter level (dist_join.cu) --
 * One process per GPU.  Replaces partition_only's shared partitions (partitioned_hash.h:82-124)
 * at node scale.
 * setup: plan for rows_total rows per relation over `world` GPUs (power of two, <= 8), allocate
 *   this rank's level-1 regions and signal words, export their CUDA IPC handles
 *   (hmj_dist_handle_bytes() bytes).
 * connect: all ranks' handles in rank order (world x hmj_dist_handle_bytes() bytes).
 * join_u64: ONE call per join on every rank (same order on all ranks), no host barrier around it:
 *   level 1 over the rank's rows of R, then of S, cut on the GPU bits and written into the owners'
 *   memory over NVLink; the owners' local levels of R run while S is still in flight; then the
 *   bucket join of the rank's own hash range.  The GPUs meet through release/acquire flags in each
 *   other's memory.  Row ids are global: r_rowid_base + i / s_rowid_base + i.  *ok = 0 when a
 *   region overflowed here or a peer did not answer (every rank then uses the NCCL path).
 * last_exchange: pairs this rank wrote into other ranks' memory in the last join and the pairs of
 *   R and S it received, read back from the device cursors.
 * teardown: every rank must have finished its last join (a barrier) before any rank tears down. */
int hmj_dist_handle_bytes(void);
int hmj_dist_setup(hmj_ctx* ctx, int rank, int world, size_t rows_total, unsigned char* handles_out);
int hmj_dist_connect(hmj_ctx* ctx, const unsigned char* all_handles);
int hmj_dist_join_u64(hmj_ctx* ctx, const uint64_t* d_rkeys, size_t nr, uint32_t r_rowid_base,
                      const uint64_t* d_skeys, size_t ns, uint32_t s_rowid_base,
                      uint64_t* h_num_pairs, int* ok);
int hmj_dist_last_exchange(hmj_ctx* ctx, uint64_t* sent_pairs_remote, uint64_t* received_r,
                           uint64_t* received_s);
int hmj_dist_teardown(hmj_ctx* ctx);

/* ---- string keys over several GPUs: the key gather fused with the exchange (dist_strings.cu) --
 * The rows of a relation travel to the GPU owning their hash range WITH their keys
 * (partition_only's row movement, partitioned_hash.h:102-110).  setup allocates this rank's
 * receive buffers for both relations (fixed capacities) and exports 6 IPC handles; connect takes
 * all ranks' handles in rank order.  send: the rank's rows in send order (grouped by owner) --
 * d_start/d_send_off from hmj_string_row_spans + a prefix sum -- are written straight into the
 * owners' buffers over NVLink by ONE kernel (key bytes, row id, length); the h_* arrays say which
 * send-order rows go to which owner and where this rank's block starts there (from the G x G count
 * matrix).  The caller orders the ranks (stream-ordered barrier after the sends, and before the
 * buffers are written again).  buffers: this rank's receive buffers. */
int hmj_xchg_handle_bytes(void);
int hmj_xchg_setup(hmj_ctx* ctx, int rank, int world, size_t rows_cap, size_t bytes_cap,
                   unsigned char* handles_out);
int hmj_xchg_connect(hmj_ctx* ctx, const unsigned char* all_handles);
int hmj_xchg_send(hmj_ctx* ctx, int side, const uint8_t* d_chars, const uint64_t* d_start,
                  const uint64_t* d_send_off, const uint32_t* d_rowid, const uint32_t* d_lens, size_t n,
                  const uint64_t* h_part_first, const uint64_t* h_row_base, const uint64_t* h_byte_base);
int hmj_xchg_buffers(hmj_ctx* ctx, int side, uint8_t** d_chars, uint32_t** d_rowid, uint32_t** d_lens,
                     size_t* rows_cap, size_t* bytes_cap);
int hmj_xchg_teardown(hmj_ctx* ctx);

/* ---- benchmark relation generators on the device (datagen.cu; harness, not on the join path) --
 * The reference builds its inputs on the host with create_strvec (strgen.cc:28-61: ceil(sqrt(n))
 * base words w_i, then "w_i-w_j" row-major, shuffled); these produce the same shapes in HBM.
 * permute: d_out[i] = P(d_in[i]) (d_in NULL: P(in_base + i)), P a seeded bijection of [0, domain).
 * zipf: d_out_rank[i] = first k with d_cdf[k] > u(seed, row_base + i), u uniform in [0,1).
 * strgen_lengths / strgen_write: length and bytes of strgen key d_key_ids[i] (id < q: w_id, else
 * w_a-w_b with (a, b) = divmod(id - q, q)) over a device word table (offsets[q+1] + chars). */
int hmj_gen_permute(hmj_ctx* ctx, const uint32_t* d_in, uint32_t in_base, size_t n, uint32_t domain,
                    uint32_t seed, uint32_t* d_out);
int hmj_gen_zipf(hmj_ctx* ctx, const double* d_cdf, uint32_t n_keys, uint64_t seed, uint64_t row_base,
                 size_t n, uint32_t* d_out_rank);
int hmj_gen_strgen_lengths(hmj_ctx* ctx, const uint64_t* d_word_offsets, uint32_t q,
                           const uint32_t* d_key_ids, size_t n, uint64_t* d_out_len);
int hmj_gen_strgen_write(hmj_ctx* ctx, const uint64_t* d_word_offsets, const uint8_t* d_word_chars,
                         uint32_t q, const uint32_t* d_key_ids, size_t n, const uint64_t* d_offsets,
                         uint8_t* d_chars);

/* ---- per-stage device timings of the last whole-operator call ------------------------ */
typedef struct hmj_timings {
  float hash_ms;       /* K1 */
  float histogram_ms;  /* K2 */
  float sort_ms;       /* K3 (all passes, both relations) */
  float join_ms;       /* K4-K6 */
  float h2d_ms, d2h_ms;
  float total_ms;
  uint32_t sort_passes; /* scatter passes actually executed (both relations) */
  /* dominant kernel (onesweep pass) timed per launch while profiling is on */
  float pass_ms_sum;     /* sum of the per-launch durations since profiling was (re)enabled */
  uint32_t pass_launches;
  uint32_t sort_finishes;  /* bucket-finish launches (prefix sort + small-bucket finish) */
  uint32_t sort_fallbacks; /* sorts that met an oversized bucket and re-ran with scatter passes */
  uint32_t sort_uniform;   /* sorts completed by the uniform-hash path (bucket scatter + finish) */
  /* per kernel class (hmj_kernel_class), accumulated while profiling is on */
  uint32_t kernel_launches[8];
  float kernel_ms[8];
  double kernel_bytes[8]; /* algorithmic bytes of the timed launches */
  /* plan of the last whole-operator join */
  uint32_t bucket_join;    /* 1: radix-partition + in-shared-memory bucket join; 0: sort + merge-path join */
  uint32_t scatter_levels; /* bucket-scatter launches of that join (both relations) */
} hmj_timings;
enum hmj_kernel_class {
  HMJ_K_PASS = 0,    /* onesweep_pass_kernel (stable scatter pass) */
  HMJ_K_SCATTER = 1, /* bucket_scatter_kernel (uniform-hash levels) */
  HMJ_K_FINISH = 2,  /* segment_sort_kernel / bucket_finish_kernel */
  HMJ_K_JOIN = 3,    /* join_tile_kernel */
  HMJ_K_HIST = 4,    /* radix_histogram_kernel */
  HMJ_K_HASH = 5,    /* hash_strings_kernel */
  HMJ_K_BJOIN = 6,   /* bucket_join_kernel (uniform-hash whole-operator path) */
  HMJ_K_XSCATTER = 7 /* level 1 of the multi-GPU join: bucket scatter into the owners' memory over NVLink */
};
int hmj_get_timings(hmj_ctx* ctx, hmj_timings* out);
/* on != 0: bracket every onesweep-pass launch with CUDA events on the context's stream and
 * reset the accumulated figures; 0: stop.  Used by bench.py for the roofline of the dominant
 * kernel. */
int hmj_set_profiling(hmj_ctx* ctx, int on);

/* ---- test aid (host only, needs no GPU) ------------------------------------------------
 * The bucket plan the join derives from a sample of both relations' hashes (bucket_join.cu,
 * hmj_plan_from_sample; the counterpart of the reference's choice of the partition width,
 * radix_hash.h:38-57, made from the data instead of from the row count alone).  h_sample:
 * per_side evenly spaced hashes of R, then per_side of S.  out10 = levels, bits[0], bits[1],
 * total bucket bits, skipped prefix bits, slots per bucket, prefix checked (0/1), prefix,
 * density x 1e6, skew verdict (bit 0: R, bit 1: S). */
int hmj_debug_plan_from_sample(const uint64_t* h_sample, uint32_t per_side, size_t nr, size_t ns,
                               int64_t* out10);

#ifdef __cplusplus
}
#endif
#endif /* HMJ_H */
