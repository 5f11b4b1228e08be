// radix_hash.h (facade) -- same names, template parameters and argument order as
// dryman/HashMergeJoin radix_hash.h, backed by the B200 hash + radix sort (libhmj.so).
//
//   radix_hash::optimal_partition(n)                                  reference :38-57
//   radix_hash::radix_non_inplace_par<Key,Value[,Hash]>(begin,end,dst,num_threads[,bits])  :351-423
//   radix_hash::radix_inplace_seq<Key,Value>(dst,n[,bits])            :425-493
//   radix_hash::radix_inplace_par(dst,n,num_threads[,bits])           :589-654
//
// Input items are pair-like (`->first` key, `->second` value, reference :343-344); dst items are
// std::tuple<std::size_t, Key, Value>.  partition_bits is accepted and ignored (the digit width is
// the kernel's); num_threads is the number of HOST threads used for what stays on the host
// (flattening keys, writing the sorted tuples) -- the sort itself runs on the GPU.  Result: hash column ascending
// (what the reference's tests pin); inside an equal-hash run rows keep their input order (the
// reference leaves that order unspecified).
#ifndef RADIX_HASH_H
#define RADIX_HASH_H 1

#include <tuple>
#include <utility>

#include "hmj_facade.h"

namespace radix_hash {

inline int optimal_partition(std::size_t input_num) { return hmj_optimal_partition(input_num); }

namespace detail {
struct first_of {
  template <typename Item>
  auto operator()(const Item& it) const -> decltype((it.first)) { return it.first; }
};
struct tuple_key {
  template <typename Item>
  auto operator()(const Item& it) const -> decltype((std::get<1>(it))) { return std::get<1>(it); }
};
}  // namespace detail

template <typename Key, typename Value, typename Hash = std::hash<Key>,
          typename BidirectionalIterator, typename RandomAccessIterator>
void radix_non_inplace_par(BidirectionalIterator begin, BidirectionalIterator end,
                           RandomAccessIterator dst, int num_threads, int /*partition_bits*/) {
  // num_threads: the host threads that flatten the keys and write the sorted tuples (the sort
  // itself runs on the GPU)
  const int threads = num_threads > 0 ? num_threads : 1;
  const hmj::detail::KeyColumns cols =
      hmj::detail::flatten<Key, Hash>(begin, end, detail::first_of(), threads);
  hmj::detail::PinnedArray<uint64_t> hash;
  hmj::detail::PinnedArray<uint32_t> rowid;
  hmj::detail::hash_sort<Key, Hash>(cols, &hash, &rowid);
  // row id -> element: bidirectional iterators are walked once into a pointer table
  typedef typename std::iterator_traits<BidirectionalIterator>::value_type Item;
  std::vector<const Item*> rows;
  rows.reserve(cols.n);
  for (BidirectionalIterator it = begin; it != end; ++it) rows.push_back(&*it);
  hmj::detail::parallel_slices(cols.n, threads, [&](int, std::size_t lo, std::size_t hi) {
    for (std::size_t i = lo; i < hi; ++i) {
      const Item& src = *rows[rowid[i]];
      dst[i] = std::tuple<std::size_t, Key, Value>(static_cast<std::size_t>(hash[i]), src.first,
                                                  src.second);
    }
  });
}

template <typename Key, typename Value, typename Hash = std::hash<Key>,
          typename BidirectionalIterator, typename RandomAccessIterator>
void radix_non_inplace_par(BidirectionalIterator begin, BidirectionalIterator end,
                           RandomAccessIterator dst, int num_threads) {
  radix_non_inplace_par<Key, Value, Hash, BidirectionalIterator, RandomAccessIterator>(
      begin, end, dst, num_threads, optimal_partition(std::distance(begin, end)));
}

// Pre-hashed tuples (hash, key, value), sorted in place on the hash column.
template <typename Key, typename Value, typename RandomAccessIterator>
void radix_inplace_seq(RandomAccessIterator dst, std::size_t input_num, int /*partition_bits*/) {
  if (input_num == 0) return;
  std::vector<uint64_t> hash(input_num), sorted(input_num);
  std::vector<uint32_t> rowid(input_num);
  for (std::size_t i = 0; i < input_num; ++i) hash[i] = static_cast<uint64_t>(std::get<0>(dst[i]));
  hmj::Context::check(hmj_sort_pairs_host(hmj::Context::get(), hash.data(), nullptr, input_num,
                                          sorted.data(), rowid.data()), "hmj_sort_pairs_host");
  hmj::detail::permute_in_place(dst, rowid);
}

template <typename Key, typename Value, typename RandomAccessIterator>
void radix_inplace_seq(RandomAccessIterator dst, std::size_t input_num) {
  radix_inplace_seq<Key, Value, RandomAccessIterator>(dst, input_num, optimal_partition(input_num));
}

template <typename RandomAccessIterator>
void radix_inplace_par(RandomAccessIterator dst, std::size_t input_num, int /*num_threads*/,
                       int partition_bits) {
  typedef typename std::tuple_element<1, typename RandomAccessIterator::value_type>::type Key;
  typedef typename std::tuple_element<2, typename RandomAccessIterator::value_type>::type Value;
  radix_inplace_seq<Key, Value, RandomAccessIterator>(dst, input_num, partition_bits);
}

template <typename RandomAccessIterator>
void radix_inplace_par(RandomAccessIterator dst, std::size_t input_num, int num_threads) {
  radix_inplace_par(dst, input_num, num_threads, optimal_partition(input_num));
}

}  // namespace radix_hash

#endif  // RADIX_HASH_H
