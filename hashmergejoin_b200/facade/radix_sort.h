// radix_sort.h (facade) -- ::radix_int_inplace / ::radix_int_non_inplace of
// dryman/HashMergeJoin radix_sort.h:333-398,452-522: sort of std::pair<UnsignedKey, Value> on the
// key, no hashing.  Same GPU sort kernels as radix_hash.h with the hash stage skipped.
#ifndef RADIX_SORT_H
#define RADIX_SORT_H 1

#include <utility>

#include "radix_hash.h"

template <typename Key, typename Value, typename RandomAccessIterator>
void radix_int_inplace(RandomAccessIterator dst, std::size_t input_num, int /*num_threads*/,
                       int /*partition_bits*/) {
  static_assert(std::is_unsigned<Key>::value, "Key must be an unsigned arithmic type.");
  if (input_num == 0) return;
  std::vector<uint64_t> keys(input_num), sorted(input_num);
  std::vector<uint32_t> rowid(input_num);
  for (std::size_t i = 0; i < input_num; ++i) keys[i] = static_cast<uint64_t>(dst[i].first);
  hmj::Context::check(hmj_sort_pairs_host(hmj::Context::get(), keys.data(), nullptr, input_num,
                                          sorted.data(), rowid.data()), "hmj_sort_pairs_host");
  hmj::detail::permute_in_place(dst, rowid);
}

template <typename Key, typename Value, typename RandomAccessIterator>
void radix_int_inplace(RandomAccessIterator dst, std::size_t input_num, int num_threads) {
  radix_int_inplace<Key, Value, RandomAccessIterator>(dst, input_num, num_threads,
                                                      radix_hash::optimal_partition(input_num));
}

template <typename Key, typename Value, typename BidirectionalIterator, typename RandomAccessIterator>
void radix_int_non_inplace(BidirectionalIterator begin, BidirectionalIterator end,
                           RandomAccessIterator dst, int /*num_threads*/, int /*partition_bits*/) {
  static_assert(std::is_unsigned<Key>::value, "Key must be an unsigned arithmic type.");
  typedef typename std::iterator_traits<BidirectionalIterator>::value_type Item;
  std::vector<const Item*> rows;
  std::vector<uint64_t> keys;
  for (BidirectionalIterator it = begin; it != end; ++it) {
    rows.push_back(&*it);
    keys.push_back(static_cast<uint64_t>(it->first));
  }
  const std::size_t n = keys.size();
  if (n == 0) return;
  std::vector<uint64_t> sorted(n);
  std::vector<uint32_t> rowid(n);
  hmj::Context::check(hmj_sort_pairs_host(hmj::Context::get(), keys.data(), nullptr, n, sorted.data(),
                                          rowid.data()), "hmj_sort_pairs_host");
  for (std::size_t i = 0; i < n; ++i) dst[i] = *rows[rowid[i]];
}

template <typename Key, typename Value, typename BidirectionalIterator, typename RandomAccessIterator>
void radix_int_non_inplace(BidirectionalIterator begin, BidirectionalIterator end,
                           RandomAccessIterator dst, int num_threads) {
  radix_int_non_inplace<Key, Value, BidirectionalIterator, RandomAccessIterator>(
      begin, end, dst, num_threads, radix_hash::optimal_partition(std::distance(begin, end)));
}

#endif  // RADIX_SORT_H
