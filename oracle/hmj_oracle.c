/*
 * hmj_oracle.c -- TEST INFRASTRUCTURE ONLY.  See hmj_oracle.h for scope and pinning status.
 * Every function cites the reference lines it restates; arrays are flat (hash[], rowid[])
 * where the reference uses std::tuple<size_t,Key,Value> / std::pair<Key,Value>.
 */
#include "hmj_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ std::hash<std::string>
 * libstdc++ 13.3 hash_bytes.cc, 64-bit variant (not vendored in the reference; the reference
 * reaches it via std::hash<Key> at radix_hash.h:314,340,353 and partitioned_hash.h:40,58,76). */
static inline uint64_t shift_mix(uint64_t v) { return v ^ (v >> 47); }
static inline uint64_t load_le64(const uint8_t* p) {
  uint64_t r;
  memcpy(&r, p, 8);
  return r; /* x86-64: little endian */
}
static inline uint64_t load_tail(const uint8_t* p, int n) {
  uint64_t r = 0;
  for (--n; n >= 0; --n) r = (r << 8) + p[n];
  return r;
}
uint64_t oracle_hash_bytes(const void* ptr, size_t len, uint64_t seed) {
  static const uint64_t mul = (((uint64_t)0xc6a4a793UL) << 32UL) + (uint64_t)0x5bd1e995UL;
  const uint8_t* buf = (const uint8_t*)ptr;
  const size_t len_aligned = len & ~(size_t)7;
  const uint8_t* end = buf + len_aligned;
  uint64_t hash = seed ^ (len * mul);
  for (const uint8_t* p = buf; p != end; p += 8) {
    const uint64_t data = shift_mix(load_le64(p) * mul) * mul;
    hash ^= data;
    hash *= mul;
  }
  if ((len & 7) != 0) {
    const uint64_t data = load_tail(end, (int)(len & 7));
    hash ^= data;
    hash *= mul;
  }
  hash = shift_mix(hash) * mul;
  hash = shift_mix(hash);
  return hash;
}
void oracle_hash_strings(const uint64_t* offsets, const uint8_t* chars, size_t n, uint64_t* out) {
  for (size_t i = 0; i < n; ++i)
    out[i] = oracle_hash_bytes(chars + offsets[i], (size_t)(offsets[i + 1] - offsets[i]),
                               ORACLE_STD_HASH_SEED);
}

/* ------------------------------------------------------------------ radix_hash.h:38-57 */
int oracle_optimal_partition(size_t input_num) {
  double min_dist = 1.0;
  int candidate = 0;
  for (int k = 6; k < 15; k++) {
    const double lg = log((double)input_num) / log((double)(1 << k));
    const double up = ceil(lg) - lg, down = lg - floor(lg);
    const double dist = down < up ? down : up;
    if (input_num < ((size_t)1 << k)) return k;
    if (dist <= min_dist) {
      candidate = k;
      min_dist = dist;
    }
  }
  return candidate;
}

/* ------------------------------------------------------------------ key comparisons
 * std::string operator< / operator== (lexicographic on unsigned bytes, then length) and
 * integer < / == ; used where the reference compares std::get<1> of two tuples. */
static int key_cmp(const oracle_keys* a, uint32_t ra, const oracle_keys* b, uint32_t rb) {
  if (a->kind == ORACLE_KEY_U64) {
    const uint64_t x = a->u64[ra], y = b->u64[rb];
    return x < y ? -1 : (x > y ? 1 : 0);
  }
  const size_t la = (size_t)(a->offsets[ra + 1] - a->offsets[ra]);
  const size_t lb = (size_t)(b->offsets[rb + 1] - b->offsets[rb]);
  const int c = memcmp(a->chars + a->offsets[ra], b->chars + b->offsets[rb], la < lb ? la : lb);
  if (c) return c < 0 ? -1 : 1;
  return la < lb ? -1 : (la > lb ? 1 : 0);
}

typedef struct {
  uint64_t* h;
  uint32_t* r;
  const oracle_keys* keys;
} cols;

static inline void swap_elem(cols* c, size_t i, size_t j) {
  const uint64_t th = c->h[i];
  c->h[i] = c->h[j];
  c->h[j] = th;
  const uint32_t tr = c->r[i];
  c->r[i] = c->r[j];
  c->r[j] = tr;
}

/* bf6_insertion_inner / bf6_insertion_outer, radix_hash.h:86-120: insertion sort by hash, ties
 * by key. */
static void insertion_sort(cols* c, size_t begin, size_t end) {
  for (size_t idx0 = begin + 1; idx0 < end; idx0++) {
    size_t idx = idx0;
    while (idx > begin) {
      const uint64_t h1 = c->h[idx], h2 = c->h[idx - 1];
      if (h1 > h2) break;
      if (h1 < h2 || key_cmp(c->keys, c->r[idx], c->keys, c->r[idx - 1]) < 0) {
        swap_elem(c, idx, idx - 1);
        idx--;
        continue;
      }
      break;
    }
  }
}

/* The body shared by bf6_helper_p (radix_hash.h:202-292) and bf6_helper_s (:122-200): for each
 * super bucket: skip (<2), insertion sort (< 2^(bits/2)), else count the next digit, permute in
 * place by cycle leaders, and recurse while mask bits remain.  first[]/second[] are the
 * (begin,end) pairs of `super_indexes`. */
static void msd_recurse(cols* c, const size_t* sup_first, const size_t* sup_second, int mask_bits,
                        int partition_bits) {
  const int partitions = 1 << partition_bits;
  const size_t sqrt_partitions = (size_t)1 << (partition_bits / 2);
  const uint64_t mask = mask_bits >= 64 ? ~0ULL : ((1ULL << mask_bits) - 1ULL);
  const int shift = mask_bits < partition_bits ? 0 : mask_bits - partition_bits;
  size_t* counters = (size_t*)malloc(sizeof(size_t) * partitions);
  size_t* first = (size_t*)malloc(sizeof(size_t) * partitions);
  size_t* second = (size_t*)malloc(sizeof(size_t) * partitions);
  for (int s = 0; s < partitions; s++) {
    const size_t s_begin = sup_first[s], s_end = sup_second[s];
    if (s_end - s_begin < 2) continue;
    if (s_end - s_begin < sqrt_partitions) {
      insertion_sort(c, s_begin, s_end);
      continue;
    }
    for (int i = 0; i < partitions; i++) counters[i] = 0;
    first[0] = s_begin;
    for (size_t i = s_begin; i < s_end; i++) counters[(c->h[i] & mask) >> shift]++;
    for (int i = 0; i < partitions - 1; i++) second[i] = first[i + 1] = first[i] + counters[i];
    second[partitions - 1] = first[partitions - 1] + counters[partitions - 1];
    int iter = 0;
    while (iter < partitions) {
      const size_t idx_i = first[iter];
      if (idx_i >= second[iter]) {
        iter++;
        continue;
      }
      int idx_c = (int)((c->h[idx_i] & mask) >> shift);
      if (idx_c == iter) {
        first[iter]++;
        continue;
      }
      /* tmp_bucket = move(dst[idx_i]); do { swap(dst[idx_j], tmp) } while (idx_j > idx_i) */
      uint64_t th = c->h[idx_i];
      uint32_t tr = c->r[idx_i];
      size_t idx_j;
      do {
        idx_c = (int)((th & mask) >> shift);
        idx_j = first[idx_c]++;
        const uint64_t xh = c->h[idx_j];
        const uint32_t xr = c->r[idx_j];
        c->h[idx_j] = th;
        c->r[idx_j] = tr;
        th = xh;
        tr = xr;
      } while (idx_j > idx_i);
    }
    const int new_mask_bits = mask_bits - partition_bits;
    if (new_mask_bits <= 0) continue;
    first[0] = s_begin;
    for (int i = 1; i < partitions; i++) first[i] = second[i - 1];
    msd_recurse(c, first, second, new_mask_bits, partition_bits);
  }
  free(counters);
  free(first);
  free(second);
}

/* radix_non_inplace_par, radix_hash.h:351-423.  The worker (:294-346) counts the top `bits`
 * bits per thread slice, the leader prefix-sums partition-major / thread-minor, and each thread
 * scatters its slice in input order: for any thread count that is a STABLE counting scatter on
 * h >> (64-bits), which is what is restated here.  Then bf6_helper_p over the top buckets. */
void oracle_radix_sort_non_inplace(uint64_t* hash, uint32_t* rowid, size_t n, int bits,
                                   const oracle_keys* keys) {
  if (bits <= 0) bits = oracle_optimal_partition(n);
  const int partitions = 1 << bits;
  const int shift = 64 - bits;
  size_t* first = (size_t*)calloc(partitions, sizeof(size_t));
  size_t* second = (size_t*)calloc(partitions, sizeof(size_t));
  size_t* cursor = (size_t*)calloc(partitions, sizeof(size_t));
  uint64_t* dh = (uint64_t*)malloc(sizeof(uint64_t) * (n ? n : 1));
  uint32_t* dr = (uint32_t*)malloc(sizeof(uint32_t) * (n ? n : 1));
  for (size_t i = 0; i < n; i++) cursor[hash[i] >> shift]++;
  size_t acc = 0;
  for (int p = 0; p < partitions; p++) {
    first[p] = acc;
    acc += cursor[p];
    second[p] = acc;
    cursor[p] = first[p];
  }
  for (size_t i = 0; i < n; i++) {
    const size_t d = cursor[hash[i] >> shift]++;
    dh[d] = hash[i];
    dr[d] = rowid[i];
  }
  memcpy(hash, dh, sizeof(uint64_t) * n);
  memcpy(rowid, dr, sizeof(uint32_t) * n);
  free(dh);
  free(dr);
  free(cursor);
  cols c = {hash, rowid, keys};
  msd_recurse(&c, first, second, 64 - bits, bits);
  free(first);
  free(second);
}

/* radix_inplace_seq, radix_hash.h:425-483: count top bits, cycle-leader permutation in place
 * (no "already home" shortcut at this level), then bf6_helper_p. */
void oracle_radix_sort_inplace_seq(uint64_t* hash, uint32_t* rowid, size_t n, int bits,
                                   const oracle_keys* keys) {
  if (bits <= 0) bits = oracle_optimal_partition(n);
  const int partitions = 1 << bits;
  const int shift = 64 - bits;
  size_t* counters = (size_t*)calloc(partitions, sizeof(size_t));
  size_t* first = (size_t*)calloc(partitions, sizeof(size_t));
  size_t* second = (size_t*)calloc(partitions, sizeof(size_t));
  for (size_t i = 0; i < n; i++) counters[hash[i] >> shift]++;
  first[0] = 0;
  for (int i = 0; i < partitions - 1; i++) second[i] = first[i + 1] = first[i] + counters[i];
  second[partitions - 1] = first[partitions - 1] + counters[partitions - 1];
  int iter = 0;
  while (iter < partitions) {
    const size_t idx_i = first[iter];
    if (idx_i >= second[iter]) {
      iter++;
      continue;
    }
    uint64_t th = hash[idx_i];
    uint32_t tr = rowid[idx_i];
    size_t idx_j;
    do {
      const int idx_c = (int)(th >> shift);
      idx_j = first[idx_c]++;
      const uint64_t xh = hash[idx_j];
      const uint32_t xr = rowid[idx_j];
      hash[idx_j] = th;
      rowid[idx_j] = tr;
      th = xh;
      tr = xr;
    } while (idx_j > idx_i);
  }
  first[0] = 0;
  for (int i = 1; i < partitions; i++) first[i] = second[i - 1];
  cols c = {hash, rowid, keys};
  msd_recurse(&c, first, second, 64 - bits, bits);
  free(counters);
  free(first);
  free(second);
}

/* ------------------------------------------------------------------ canonical orders */
typedef struct {
  uint64_t h;
  uint32_t r;
} hr_pair;
static int cmp_hr(const void* a, const void* b) {
  const hr_pair *x = (const hr_pair*)a, *y = (const hr_pair*)b;
  if (x->h != y->h) return x->h < y->h ? -1 : 1;
  return x->r < y->r ? -1 : (x->r > y->r ? 1 : 0);
}
void oracle_canonical_sort(uint64_t* hash, uint32_t* rowid, size_t n) {
  hr_pair* v = (hr_pair*)malloc(sizeof(hr_pair) * (n ? n : 1));
  for (size_t i = 0; i < n; i++) {
    v[i].h = hash[i];
    v[i].r = rowid[i];
  }
  qsort(v, n, sizeof(hr_pair), cmp_hr);
  for (size_t i = 0; i < n; i++) {
    hash[i] = v[i].h;
    rowid[i] = v[i].r;
  }
  free(v);
}
static int cmp_u64(const void* a, const void* b) {
  const uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
  return x < y ? -1 : (x > y ? 1 : 0);
}
void oracle_sort_pairs(uint32_t* out_r, uint32_t* out_s, size_t m) {
  uint64_t* v = (uint64_t*)malloc(sizeof(uint64_t) * (m ? m : 1));
  for (size_t i = 0; i < m; i++) v[i] = ((uint64_t)out_r[i] << 32) | out_s[i];
  qsort(v, m, sizeof(uint64_t), cmp_u64);
  for (size_t i = 0; i < m; i++) {
    out_r[i] = (uint32_t)(v[i] >> 32);
    out_s[i] = (uint32_t)v[i];
  }
  free(v);
}

/* ------------------------------------------------------------------ hashjoin.h:73-154
 * The iterator of HashMergeJoin as a loop.  find_match = the while(true) of :77-102 / :126-153;
 * the step before it is operator++ :105-124. */
size_t oracle_merge_walk_reference(const uint64_t* rh, const uint32_t* rr, size_t nr,
                                   const uint64_t* sh, const uint32_t* sr, size_t ns,
                                   const oracle_keys* rkeys, const oracle_keys* skeys,
                                   uint32_t* out_r, uint32_t* out_s, size_t cap) {
  size_t ri = 0, si = 0, m = 0;
  int first = 1;
  for (;;) {
    if (!first) {
      /* operator++ */
      if (ri + 1 == nr || si + 1 == ns) break;          /* :105-114 -> end() */
      if (rh[ri] == rh[ri + 1]) ri++;                   /* :115-118 */
      else if (sh[si] == sh[si + 1]) si++;              /* :119-122 */
      else { ri++; si++; }                              /* :123-124 */
    }
    first = 0;
    /* find_match */
    for (;;) {
      if (ri == nr || si == ns) { ri = nr; si = ns; break; }
      if (rh[ri] < sh[si]) { ri++; continue; }
      if (sh[si] < rh[ri]) { si++; continue; }
      const int c = key_cmp(rkeys, rr[ri], skeys, sr[si]);
      if (c == 0) break;
      if (c < 0) ri++; else si++;
    }
    if (ri == nr) break; /* == end() */
    if (m < cap && out_r) {
      out_r[m] = rr[ri];
      out_s[m] = sr[si];
    }
    m++;
  }
  return m;
}

/* true equi-join of two sorted streams: cross product of equal-hash runs, key-verified */
size_t oracle_equi_join(const uint64_t* rh, const uint32_t* rr, size_t nr, const uint64_t* sh,
                        const uint32_t* sr, size_t ns, const oracle_keys* rkeys,
                        const oracle_keys* skeys, uint32_t* out_r, uint32_t* out_s, size_t cap,
                        uint64_t* sum) {
  size_t i = 0, j = 0, m = 0;
  uint64_t acc = 0;
  while (i < nr && j < ns) {
    if (rh[i] < sh[j]) { i++; continue; }
    if (sh[j] < rh[i]) { j++; continue; }
    const uint64_t h = rh[i];
    size_t i1 = i, j1 = j;
    while (i1 < nr && rh[i1] == h) i1++;
    while (j1 < ns && sh[j1] == h) j1++;
    for (size_t a = i; a < i1; a++)
      for (size_t b = j; b < j1; b++)
        if (key_cmp(rkeys, rr[a], skeys, sr[b]) == 0) {
          if (m < cap && out_r) {
            out_r[m] = rr[a];
            out_s[m] = sr[b];
          }
          acc += (uint64_t)rr[a] + sr[b];
          m++;
        }
    i = i1;
    j = j1;
  }
  if (sum) *sum = acc;
  return m;
}

/* partitioned_hash.h:57-60 (and :102 shift = 64 - partition_bits) */
void oracle_partition_counts(const uint64_t* hash, size_t n, int bits, uint64_t* counts) {
  const int shift = 64 - bits;
  for (size_t p = 0; p < ((size_t)1 << bits); p++) counts[p] = 0;
  for (size_t i = 0; i < n; i++) counts[hash[i] >> shift]++;
}

/* hashjoin_bench.cc:47-53: s_map[k] = v for every S row in order (last wins), then for every R
 * row sum += r.v + s_map[r.k] (0 when absent).  Done with a sort of S by (hash,row) and a
 * binary search per R row instead of std::unordered_map; the result is order independent. */
void oracle_hash_join_raw(const uint64_t* rhash, size_t nr, const oracle_keys* rkeys,
                          const uint64_t* rvals, const uint64_t* shash, size_t ns,
                          const oracle_keys* skeys, const uint64_t* svals, uint64_t* hits,
                          uint64_t* sum) {
  hr_pair* v = (hr_pair*)malloc(sizeof(hr_pair) * (ns ? ns : 1));
  for (size_t i = 0; i < ns; i++) {
    v[i].h = shash[i];
    v[i].r = (uint32_t)i;
  }
  qsort(v, ns, sizeof(hr_pair), cmp_hr);
  uint64_t nh = 0, acc = 0;
  for (size_t i = 0; i < nr; i++) {
    acc += rvals ? rvals[i] : (uint64_t)i;
    size_t lo = 0, hi = ns;
    while (lo < hi) {
      const size_t mid = lo + (hi - lo) / 2;
      if (v[mid].h < rhash[i]) lo = mid + 1; else hi = mid;
    }
    long best = -1;
    for (size_t j = lo; j < ns && v[j].h == rhash[i]; j++)
      if (key_cmp(rkeys, (uint32_t)i, skeys, v[j].r) == 0) best = (long)v[j].r; /* ascending rows */
    if (best >= 0) {
      nh++;
      acc += svals ? svals[best] : (uint64_t)best;
    }
  }
  free(v);
  *hits = nh;
  *sum = acc;
}

size_t oracle_join_u64(const uint64_t* rkeys, size_t nr, const uint64_t* skeys, size_t ns,
                       uint32_t* out_r, uint32_t* out_s, size_t cap, uint64_t* sum) {
  uint64_t* rh = (uint64_t*)malloc(8 * (nr ? nr : 1));
  uint64_t* sh = (uint64_t*)malloc(8 * (ns ? ns : 1));
  uint32_t* rr = (uint32_t*)malloc(4 * (nr ? nr : 1));
  uint32_t* sr = (uint32_t*)malloc(4 * (ns ? ns : 1));
  for (size_t i = 0; i < nr; i++) { rh[i] = rkeys[i]; rr[i] = (uint32_t)i; }
  for (size_t i = 0; i < ns; i++) { sh[i] = skeys[i]; sr[i] = (uint32_t)i; }
  oracle_keys rk = {ORACLE_KEY_U64, rkeys, 0, 0}, sk = {ORACLE_KEY_U64, skeys, 0, 0};
  oracle_radix_sort_non_inplace(rh, rr, nr, 0, &rk);
  oracle_radix_sort_non_inplace(sh, sr, ns, 0, &sk);
  const size_t m = oracle_equi_join(rh, rr, nr, sh, sr, ns, &rk, &sk, out_r, out_s, cap, sum);
  free(rh); free(sh); free(rr); free(sr);
  return m;
}
