// hashjoin.h (facade) -- HashMergeJoin / HashMergeJoin2 of dryman/HashMergeJoin hashjoin.h:33-363
// with the same constructor arguments and the same input-iterator protocol
// (begin()/end()/clear(), *it -> std::tuple<Key*, RValue*, SValue*>&), backed by the B200
// hash -> radix sort -> merge join (libhmj.so).
//
// Semantics: the TRUE equi-join -- every (r, s) with equal keys, duplicate runs expanded as the
// full cross product, key equality verified on every hash match.  The reference iterator emits
// a subset on N:M runs and stops when either cursor reaches its last element (DESIGN.md,
// "reference defects"); on unique keys both agree.
// Lifetime: like the reference, dereferenced pointers alias storage owned by the operator and
// die with clear() / the operator.
#ifndef HASH_JOIN_H
#define HASH_JOIN_H 1

#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include "radix_hash.h"

typedef std::vector<std::pair<std::string, uint64_t>> KeyValVec;
typedef std::vector<std::tuple<std::size_t, std::string, uint64_t>> HashKeyValVec;

namespace hmj {
namespace detail {

// GPU join of two flattened key columns -> (build_row, probe_row) pairs, received in pinned
// arrays inside the call (the library overlaps the copies with the join, host_api.cu).
// For kHostHash keys only hash equality is decided on the device; the caller filters by key.
template <typename Key, typename Hash>
std::size_t join_columns(const KeyColumns& r, const KeyColumns& s, PinnedArray<uint32_t>* out_r,
                         PinnedArray<uint32_t>* out_s) {
  const double t0 = now_s();
  hmj_ctx* ctx = Context::get();
  uint64_t m = 0;
  const bool str = key_path<Key, Hash>::value == kString;
  // room for a join with unique keys on one side; a larger result is fetched afterwards
  const std::size_t cap = std::max(r.n, s.n);
  out_r->resize(cap);
  out_s->resize(cap);
  int rc;
  if (str)
    rc = hmj_join_strings_host(ctx, r.offsets.data(), r.chars.data(), r.n, s.offsets.data(),
                               s.chars.data(), s.n, out_r->data(), out_s->data(), cap, &m);
  else
    rc = hmj_join_u64_host(ctx, r.u64.data(), r.n, s.u64.data(), s.n, out_r->data(), out_s->data(), cap, &m);
  if (rc == HMJ_ECAPACITY && m > cap) {
    out_r->resize(m);
    out_s->resize(m);
    rc = hmj_join_result_to_host(ctx, out_r->data(), out_s->data(), m);
  }
  Context::check(rc, str ? "hmj_join_strings_host" : "hmj_join_u64_host");
  stages().device_s += now_s() - t0;
  return (std::size_t)m;
}

// The iterator both operators hand out: walks the materialised pair list.
template <typename Key, typename RValue, typename SValue>
class JoinCursor : public std::iterator<std::input_iterator_tag, std::tuple<Key*, RValue*, SValue*>> {
 public:
  typedef std::tuple<Key*, RValue*, SValue*> value_type;
  JoinCursor(value_type* rows, std::size_t pos) : rows_(rows), pos_(pos) {}
  JoinCursor& operator++() { ++pos_; return *this; }
  JoinCursor operator++(int) { JoinCursor t(*this); ++*this; return t; }
  bool operator==(const JoinCursor& o) const { return pos_ == o.pos_; }
  bool operator!=(const JoinCursor& o) const { return pos_ != o.pos_; }
  value_type& operator*() { return rows_[pos_]; }
  value_type* operator->() { return rows_ + pos_; }

 private:
  value_type* rows_;
  std::size_t pos_;
};

}  // namespace detail
}  // namespace hmj

template <typename RIter, typename SIter>
class HashMergeJoin {
  static_assert(std::is_same<typename RIter::value_type::first_type,
                             typename SIter::value_type::first_type>::value,
                "RIter and SIter key type must be the same");
  typedef typename RIter::value_type::first_type Key;
  typedef typename RIter::value_type::second_type RValue;
  typedef typename SIter::value_type::second_type SValue;
  typedef std::hash<Key> Hash;

 public:
  typedef hmj::detail::JoinCursor<Key, RValue, SValue> iterator;

  HashMergeJoin() = default;
  HashMergeJoin(RIter r_begin, RIter r_end, SIter s_begin, SIter s_end,
                unsigned int num_threads = 1) {
    namespace d = hmj::detail;
    d::stages() = d::Stages{0, 0, 0, 0};
    const int threads = (int)(num_threads ? num_threads : 1);
    // The operator owns copies of both relations, as the reference's does (_r_sorted/_s_sorted,
    // :62-63): dereferenced pointers stay valid when the caller's containers go away.  Callers
    // whose inputs outlive the operator can compile with -DHMJ_FACADE_BORROW_INPUTS and skip it
    // (random-access inputs only; for std::string keys the copy is most of the host time).
    double t0 = d::now_s();
    copy_range(r_begin, r_end, &_r, &_r_data, threads);
    copy_range(s_begin, s_end, &_s, &_s_data, threads);
    d::stages().copy_s = d::now_s() - t0;
    const std::size_t nr = (std::size_t)std::distance(r_begin, r_end), ns = (std::size_t)std::distance(s_begin, s_end);
    const d::KeyColumns rc = d::flatten<Key, Hash>(_r_data, _r_data + nr, radix_hash::detail::first_of(), threads);
    const d::KeyColumns sc = d::flatten<Key, Hash>(_s_data, _s_data + ns, radix_hash::detail::first_of(), threads);
    d::PinnedArray<uint32_t> pr, ps;
    const std::size_t m = d::join_columns<Key, Hash>(rc, sc, &pr, &ps);
    t0 = d::now_s();
    const bool verify = d::key_path<Key, Hash>::value == d::kHostHash;
    std::pair<Key, RValue>* R = _r_data;
    std::pair<Key, SValue>* S = _s_data;
    if (verify) {
      // hashes of a user functor can collide: keep the pairs whose keys are equal
      std::vector<uint32_t> keep;
      for (std::size_t i = 0; i < m; ++i)
        if (R[pr[i]].first == S[ps[i]].first) keep.push_back((uint32_t)i);
      _rows.build(keep.size(), threads, [&](std::size_t k) {
        const std::size_t i = keep[k];
        return std::make_tuple(&R[pr[i]].first, &R[pr[i]].second, &S[ps[i]].second);
      });
    } else {
      _rows.build(m, threads, [&](std::size_t i) {
        return std::make_tuple(&R[pr[i]].first, &R[pr[i]].second, &S[ps[i]].second);
      });
    }
    d::stages().rows_s = d::now_s() - t0;
  }

  iterator begin() { return iterator(_rows.begin(), 0); }
  iterator end() { return iterator(_rows.begin(), _rows.size()); }
  void clear() {
    _rows.clear();
    _r.clear();
    _s.clear();
    _r_data = nullptr;
    _s_data = nullptr;
  }

 protected:
  template <typename Iter, typename Item>
  static void copy_range(Iter b, Iter e, hmj::detail::ParallelArray<Item>* own, Item** data, int threads) {
    const std::size_t n = (std::size_t)std::distance(b, e);
#ifdef HMJ_FACADE_BORROW_INPUTS
    if (hmj::detail::is_random_access<Iter>::value && n) {
      *data = const_cast<Item*>(&*b);
      return;
    }
#endif
    if (hmj::detail::is_random_access<Iter>::value) {
      own->build(n, threads, [&](std::size_t i) { return *hmj::detail::advance_to(b, i); });
    } else {
      std::vector<const Item*> at;
      at.reserve(n);
      for (Iter it = b; it != e; ++it) at.push_back(&*it);
      own->build(n, threads, [&](std::size_t i) { return *at[i]; });
    }
    *data = own->begin();
  }
  hmj::detail::ParallelArray<std::pair<Key, RValue>> _r;
  hmj::detail::ParallelArray<std::pair<Key, SValue>> _s;
  std::pair<Key, RValue>* _r_data = nullptr;
  std::pair<Key, SValue>* _s_data = nullptr;
  hmj::detail::ParallelArray<std::tuple<Key*, RValue*, SValue*>> _rows;
};

// Pre-hashed tuples (hash, key, value) owned by the caller; both ranges are sorted in place on
// the hash column (reference :224-236 calls radix_inplace_par) and then merge-joined.
template <typename RIter, typename SIter>
class HashMergeJoin2 {
  static_assert(std::is_same<typename std::tuple_element<1, typename RIter::value_type>::type,
                             typename std::tuple_element<1, typename SIter::value_type>::type>::value,
                "RIter and SIter key type must be the same");
  typedef typename std::tuple_element<1, typename RIter::value_type>::type Key;
  typedef typename std::tuple_element<2, typename RIter::value_type>::type RValue;
  typedef typename std::tuple_element<2, typename SIter::value_type>::type SValue;

 public:
  typedef hmj::detail::JoinCursor<Key, RValue, SValue> iterator;

  HashMergeJoin2() = default;
  HashMergeJoin2(RIter r_begin, RIter r_end, SIter s_begin, SIter s_end, unsigned int num_threads = 1)
      : _r_begin(r_begin), _r_end(r_end), _s_begin(s_begin), _s_end(s_end) {
    const std::size_t nr = std::distance(r_begin, r_end), ns = std::distance(s_begin, s_end);
    radix_hash::radix_inplace_par(_r_begin, nr, (int)num_threads);
    radix_hash::radix_inplace_par(_s_begin, ns, (int)num_threads);
    // join on the caller's hash column; keys are compared on the host for every hash match
    hmj::detail::KeyColumns rc, sc;
    rc.n = nr;
    sc.n = ns;
    rc.u64.resize(nr);
    sc.u64.resize(ns);
    std::size_t k = 0;
    for (RIter it = r_begin; it != r_end; ++it) rc.u64[k++] = (uint64_t)std::get<0>(*it);
    k = 0;
    for (SIter it = s_begin; it != s_end; ++it) sc.u64[k++] = (uint64_t)std::get<0>(*it);
    hmj::detail::PinnedArray<uint32_t> pr, ps;
    struct caller_hash {};  // any non-std::hash type selects the shipped-hash (u64) path
    const std::size_t m = hmj::detail::join_columns<Key, caller_hash>(rc, sc, &pr, &ps);
    for (std::size_t i = 0; i < m; ++i) {
      auto& r = _r_begin[pr[i]];
      auto& s = _s_begin[ps[i]];
      if (!(std::get<1>(r) == std::get<1>(s))) continue;
      _rows.push_back(std::make_tuple(&std::get<1>(r), &std::get<2>(r), &std::get<2>(s)));
    }
  }

  iterator begin() { return iterator(_rows.data(), 0); }
  iterator end() { return iterator(_rows.data(), _rows.size()); }
  void clear() { _rows.clear(); }

 protected:
  RIter _r_begin, _r_end;
  SIter _s_begin, _s_end;
  std::vector<std::tuple<Key*, RValue*, SValue*>> _rows;
};

#endif  // HASH_JOIN_H
