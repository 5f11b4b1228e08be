/*
 * hmj_oracle.h -- TEST INFRASTRUCTURE ONLY (never linked into or called from the product).
 *
 * Plain-C CPU restatement of the reference's hash -> radix-sort -> merge-join path on flat
 * columnar arrays.  Parity status: PINNED -- tests/test_oracle_vs_reference.py checks every
 * function below against the unmodified reference compiled into oracle/_ref/libhmj_ref.so
 * (all radix_hash_test.cc / radix_sort_test.cc / partitioned_hash_test.cc cases plus seeded
 * random inputs), and against the committed golden vectors in tests/golden/ that were generated
 * from that library (tests/golden/make_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.
 */
#ifndef HMJ_ORACLE_H
#define HMJ_ORACLE_H 1
#include <stddef.h>
#include <stdint.h>

#define ORACLE_KEY_U64 0
#define ORACLE_KEY_STRING 1

/* key storage indexed by row id */
typedef struct oracle_keys {
  int kind;
  const uint64_t* u64;     /* ORACLE_KEY_U64 */
  const uint64_t* offsets; /* ORACLE_KEY_STRING: n+1 offsets */
  const uint8_t* chars;
} oracle_keys;

/* libstdc++ std::_Hash_bytes (64-bit), the function behind std::hash<std::string>; the
 * reference calls it through Hash{}(key) at radix_hash.h:314,340. */
uint64_t oracle_hash_bytes(const void* ptr, size_t len, uint64_t seed);
#define ORACLE_STD_HASH_SEED 0xc70f6907ULL
void oracle_hash_strings(const uint64_t* offsets, const uint8_t* chars, size_t n, uint64_t* out);

/* radix_hash::optimal_partition, radix_hash.h:38-57 */
int oracle_optimal_partition(size_t input_num);

/* radix_non_inplace_par, radix_hash.h:351-423 (+ worker :294-346, bf6_helper_p/s :122-292,
 * insertion sort :86-120): hash[]/rowid[] hold the UNSORTED column (rowid[i] = i normally) and
 * are replaced by the reference's output order.  bits <= 0 uses optimal_partition(n). */
void oracle_radix_sort_non_inplace(uint64_t* hash, uint32_t* rowid, size_t n, int bits,
                                   const oracle_keys* keys);
/* radix_inplace_seq, radix_hash.h:425-493 (unstable top-level permutation, then the same
 * recursion). */
void oracle_radix_sort_inplace_seq(uint64_t* hash, uint32_t* rowid, size_t n, int bits,
                                   const oracle_keys* keys);
/* canonical order used for comparisons: ascending (hash, rowid) */
void oracle_canonical_sort(uint64_t* hash, uint32_t* rowid, size_t n);

/* HashMergeJoin::iterator walk, hashjoin.h:73-154, on already sorted columns.  Emits what the
 * reference emits, defects included (stops at the last element of either side; N:M runs give
 * N+M-1 pairs).  Returns the number of pairs (only the first cap are stored). */
size_t oracle_merge_walk_reference(const uint64_t* rh, const uint32_t* rr, size_t nr,
                                   const uint64_t* sh, const uint32_t* sr, size_t ns,
                                   const oracle_keys* rkeys, const oracle_keys* skeys,
                                   uint32_t* out_r, uint32_t* out_s, size_t cap);
/* the true equi-join of the two sorted streams: all (r,s) with equal hash and equal key,
 * emitted in (r position, s position) order.  *sum = sum of (r_row + s_row) over pairs. */
size_t oracle_equi_join(const uint64_t* rh, const uint32_t* rr, size_t nr, const uint64_t* sh,
                        const uint32_t* sr, size_t ns, const oracle_keys* rkeys,
                        const oracle_keys* skeys, uint32_t* out_r, uint32_t* out_s, size_t cap,
                        uint64_t* sum);
/* sorts pairs lexicographically by (r,s): the canonical form of a result multiset */
void oracle_sort_pairs(uint32_t* out_r, uint32_t* out_s, size_t m);

/* partition_only's counting loop, partitioned_hash.h:57-60 */
void oracle_partition_counts(const uint64_t* hash, size_t n, int bits, uint64_t* counts);

/* BM_hash_join_raw's loops, hashjoin_bench.cc:47-53, with value = row index unless vals given */
void oracle_hash_join_raw(const uint64_t* rhash, size_t nr, const oracle_keys* rkeys,
                          const uint64_t* rvals, const uint64_t* shash, size_t ns,
                          const oracle_keys* skeys, const uint64_t* svals, uint64_t* hits,
                          uint64_t* sum);

/* whole path on flat inputs (hash, sort, true join); used as the "port" CPU baseline */
size_t oracle_join_u64(const uint64_t* rkeys, size_t nr, const uint64_t* skeys, size_t ns,
                       uint32_t* out_r, uint32_t* out_s, size_t cap, uint64_t* sum);
#endif
