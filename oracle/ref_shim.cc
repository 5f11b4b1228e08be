// ref_shim.cc -- TEST INFRASTRUCTURE ONLY.  A thin C entry layer over the UNMODIFIED reference
// headers, compiled from where they lie (-I/root/reference; see Makefile target `ref`) into
// oracle/_ref/libhmj_ref.so.  Nothing here is copied from the reference: this file only
// converts flat columnar arrays into the containers the reference templates take, calls them,
// and flattens the result.  It is used (a) to pin oracle/hmj_oracle.c and (b) as the timed
// CPU baseline (`cpu_baseline.kind = "reference"`, `bench.py --impl reference`).
// Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load the resulting library.
#include <chrono>
#include <cstdint>
#include <cstring>
#include <functional>
#include <string>
#include <thread>
#include <tuple>
#include <unordered_map>
#include <utility>
#include <vector>

#include "hashjoin.h"          // /root/reference/hashjoin.h  (HashMergeJoin, HashMergeJoin2)
#include "partitioned_hash.h"  // /root/reference/partitioned_hash.h
#include "radix_hash.h"        // /root/reference/radix_hash.h
#include "radix_sort.h"        // /root/reference/radix_sort.h

namespace {

typedef std::vector<std::pair<std::string, uint64_t>> StrRel;
typedef std::vector<std::pair<uint64_t, uint64_t>> IntRel;

StrRel make_str_rel(const uint64_t* off, const uint8_t* chars, size_t n) {
  StrRel v(n);
  for (size_t i = 0; i < n; ++i) {
    v[i].first.assign(reinterpret_cast<const char*>(chars) + off[i], off[i + 1] - off[i]);
    v[i].second = i;  // Value = row index, so results can be compared as row pairs
  }
  return v;
}
IntRel make_int_rel(const uint64_t* keys, size_t n) {
  IntRel v(n);
  for (size_t i = 0; i < n; ++i) v[i] = std::make_pair(keys[i], (uint64_t)i);
  return v;
}
double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

template <typename Rel>
double sort_rel(Rel& src, int threads, int bits, uint64_t* out_hash, uint32_t* out_rowid) {
  typedef typename Rel::value_type::first_type Key;
  std::vector<std::tuple<std::size_t, Key, uint64_t>> dst(src.size());
  const double t0 = now_s();
  if (bits > 0)
    radix_hash::radix_non_inplace_par<Key, uint64_t>(src.begin(), src.end(), dst.begin(), threads, bits);
  else
    radix_hash::radix_non_inplace_par<Key, uint64_t>(src.begin(), src.end(), dst.begin(), threads);
  const double t1 = now_s();
  for (size_t i = 0; i < dst.size(); ++i) {
    if (out_hash) out_hash[i] = std::get<0>(dst[i]);
    if (out_rowid) out_rowid[i] = (uint32_t)std::get<2>(dst[i]);
  }
  return t1 - t0;
}

template <typename Rel>
double join_rel(Rel& r, Rel& s, int threads, uint32_t* out_r, uint32_t* out_s, size_t cap,
                uint64_t* num_pairs, uint64_t* sum) {
  typedef typename Rel::iterator It;
  const double t0 = now_s();
  HashMergeJoin<It, It> hmj(r.begin(), r.end(), s.begin(), s.end(), threads);
  uint64_t m = 0, acc = 0;
  for (auto t : hmj) {  // the consumer of hashjoin_bench.cc:131-133
    const uint64_t rv = *std::get<1>(t), sv = *std::get<2>(t);
    acc += rv + sv;
    if (out_r && m < cap) {
      out_r[m] = (uint32_t)rv;
      out_s[m] = (uint32_t)sv;
    }
    ++m;
  }
  const double t1 = now_s();
  *num_pairs = m;
  if (sum) *sum = acc;
  return t1 - t0;
}

// The unordered_map build + probe loop that hashjoin_bench.cc:41-53 (BM_hash_join_raw) times
// beside BM_HashMergeJoin.  That file is a google-benchmark main and cannot be built here, so its
// two loops are restated over the same containers (reserve outside the timed part, as there).
template <typename Rel>
double hash_join_raw_rel(const Rel& r, const Rel& s, uint64_t* sum_out) {
  typedef typename Rel::value_type::first_type Key;
  std::unordered_map<Key, uint64_t> s_map;
  s_map.reserve(s.size());
  const double t0 = now_s();
  uint64_t sum = 0;
  for (const auto& s_kv : s) s_map[s_kv.first] = s_kv.second;
  for (const auto& r_kv : r) sum += r_kv.second + s_map[r_kv.first];
  const double t1 = now_s();
  if (sum_out) *sum_out = sum;
  return t1 - t0;
}

}  // namespace

extern "C" {

int ref_hardware_concurrency() { return (int)std::thread::hardware_concurrency(); }
int ref_optimal_partition(size_t n) { return radix_hash::optimal_partition(n); }

// std::hash<std::string> exactly as the reference's default Hash functor computes it
uint64_t ref_std_hash_bytes(const uint8_t* p, size_t len) {
  return std::hash<std::string>{}(std::string(reinterpret_cast<const char*>(p), len));
}
void ref_hash_strings(const uint64_t* off, const uint8_t* chars, size_t n, uint64_t* out) {
  for (size_t i = 0; i < n; ++i) out[i] = ref_std_hash_bytes(chars + off[i], off[i + 1] - off[i]);
}
uint64_t ref_std_hash_u64(uint64_t k) { return std::hash<uint64_t>{}(k); }

// radix_non_inplace_par (radix_hash.h:351-423), Value = row index.  Returns seconds.
double ref_sort_u64(const uint64_t* keys, size_t n, int threads, int bits, uint64_t* out_hash,
                    uint32_t* out_rowid) {
  IntRel src = make_int_rel(keys, n);
  return sort_rel(src, threads, bits, out_hash, out_rowid);
}
double ref_sort_strings(const uint64_t* off, const uint8_t* chars, size_t n, int threads, int bits,
                        uint64_t* out_hash, uint32_t* out_rowid) {
  StrRel src = make_str_rel(off, chars, n);
  return sort_rel(src, threads, bits, out_hash, out_rowid);
}

// radix_inplace_seq / radix_inplace_par (radix_hash.h:425-654) on pre-hashed tuples.
// mode 0 = seq, 1 = par.  hash[] and rowid[] are permuted in place.
double ref_sort_inplace(uint64_t* hash, uint32_t* rowid, size_t n, int mode, int threads, int bits) {
  std::vector<std::tuple<std::size_t, uint64_t, uint64_t>> v(n);
  for (size_t i = 0; i < n; ++i) v[i] = std::make_tuple((std::size_t)hash[i], hash[i], (uint64_t)(rowid ? rowid[i] : i));
  const double t0 = now_s();
  if (mode == 0) {
    if (bits > 0) radix_hash::radix_inplace_seq<uint64_t, uint64_t>(v.begin(), n, bits);
    else radix_hash::radix_inplace_seq<uint64_t, uint64_t>(v.begin(), n);
  } else {
    if (bits > 0) radix_hash::radix_inplace_par(v.begin(), n, threads, bits);
    else radix_hash::radix_inplace_par(v.begin(), n, threads);
  }
  const double t1 = now_s();
  for (size_t i = 0; i < n; ++i) {
    hash[i] = std::get<0>(v[i]);
    if (rowid) rowid[i] = (uint32_t)std::get<2>(v[i]);
  }
  return t1 - t0;
}

// radix_int_inplace / radix_int_non_inplace (radix_sort.h:333-398,452-522). mode 0/1.
double ref_sort_int_pairs(uint64_t* keys, uint32_t* vals, size_t n, int mode, int threads) {
  std::vector<std::pair<std::size_t, uint32_t>> v(n), dst(n);
  for (size_t i = 0; i < n; ++i) v[i] = std::make_pair((std::size_t)keys[i], vals ? vals[i] : (uint32_t)i);
  const double t0 = now_s();
  if (mode == 0) ::radix_int_inplace<std::size_t, uint32_t>(v.begin(), n, threads);
  else ::radix_int_non_inplace<std::size_t, uint32_t>(v.begin(), v.end(), dst.begin(), threads);
  const double t1 = now_s();
  const auto& res = mode == 0 ? v : dst;
  for (size_t i = 0; i < n; ++i) {
    keys[i] = res[i].first;
    if (vals) vals[i] = res[i].second;
  }
  return t1 - t0;
}

// HashMergeJoin ctor + full iteration (hashjoin.h:56-68,183-195).  Returns seconds.
double ref_join_u64(const uint64_t* rk, size_t nr, const uint64_t* sk, size_t ns, int threads,
                    uint32_t* out_r, uint32_t* out_s, size_t cap, uint64_t* num_pairs, uint64_t* sum) {
  IntRel r = make_int_rel(rk, nr), s = make_int_rel(sk, ns);
  return join_rel(r, s, threads, out_r, out_s, cap, num_pairs, sum);
}
double ref_join_strings(const uint64_t* roff, const uint8_t* rchars, size_t nr, const uint64_t* soff,
                        const uint8_t* schars, size_t ns, int threads, uint32_t* out_r,
                        uint32_t* out_s, size_t cap, uint64_t* num_pairs, uint64_t* sum) {
  StrRel r = make_str_rel(roff, rchars, nr), s = make_str_rel(soff, schars, ns);
  return join_rel(r, s, threads, out_r, out_s, cap, num_pairs, sum);
}

// std::unordered_map build (S) + probe (R), hashjoin_bench.cc:41-53.  Returns seconds.
double ref_hash_join_raw_u64(const uint64_t* rk, size_t nr, const uint64_t* sk, size_t ns, uint64_t* sum) {
  IntRel r = make_int_rel(rk, nr), s = make_int_rel(sk, ns);
  return hash_join_raw_rel(r, s, sum);
}
double ref_hash_join_raw_strings(const uint64_t* roff, const uint8_t* rchars, size_t nr,
                                 const uint64_t* soff, const uint8_t* schars, size_t ns, uint64_t* sum) {
  StrRel r = make_str_rel(roff, rchars, nr), s = make_str_rel(soff, schars, ns);
  return hash_join_raw_rel(r, s, sum);
}

// partition_only (partitioned_hash.h:82-124): sizes of the produced partitions.
void ref_partition_counts_u64(const uint64_t* keys, size_t n, int threads, int bits, uint64_t* counts) {
  IntRel src = make_int_rel(keys, n);
  std::vector<std::vector<std::pair<uint64_t, uint64_t>>> dst(size_t(1) << bits);
  radix_hash::partition_only(src.begin(), src.end(), &dst, threads, bits);
  for (size_t p = 0; p < dst.size(); ++p) counts[p] = dst[p].size();
}
void ref_partition_counts_strings(const uint64_t* off, const uint8_t* chars, size_t n, int threads,
                                  int bits, uint64_t* counts) {
  StrRel src = make_str_rel(off, chars, n);
  std::vector<std::vector<std::pair<std::string, uint64_t>>> dst(size_t(1) << bits);
  radix_hash::partition_only(src.begin(), src.end(), &dst, threads, bits);
  for (size_t p = 0; p < dst.size(); ++p) counts[p] = dst[p].size();
}

}  // extern "C"
