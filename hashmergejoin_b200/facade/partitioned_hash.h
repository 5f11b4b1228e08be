// partitioned_hash.h (facade) -- partition_only / partitioned_hash_table of dryman/HashMergeJoin
// partitioned_hash.h:82-124,173-215 with the same template parameters and arguments.
// partition = Hash{}(key) >> (64 - partition_bits), exactly the reference's function (:59,102).
// Hashing and grouping run on the GPU (hash + sort of (hash,row) pairs: ascending hash order is
// partition order); the host only moves the caller's items into the destination containers.
// Order inside a partition: ascending (hash, input position) -- deterministic, where the
// reference's order depends on thread interleaving (push_back under a mutex, :77).
#ifndef PARTITIONED_HASH_H
#define PARTITIONED_HASH_H 1

#include <tuple>
#include <unordered_map>
#include <utility>
#include <vector>

#include "hmj_facade.h"

namespace hmj {
namespace detail {
struct get0_of {
  template <typename Item>
  auto operator()(const Item& it) const -> decltype((std::get<0>(it))) { return std::get<0>(it); }
};
}  // namespace detail
}  // namespace hmj

namespace radix_hash {  // the reference declares both functions in this namespace (:34)

template <typename BidirectionalIterator,
          typename ItemType = typename BidirectionalIterator::value_type,
          typename Key = typename std::tuple_element<0, ItemType>::type,
          typename Value = typename std::tuple_element<1, ItemType>::type,
          typename Hash = std::hash<Key>>
void partition_only(BidirectionalIterator begin, BidirectionalIterator end,
                    std::vector<std::vector<ItemType>>* dst, int /*num_threads*/, int partition_bits) {
  const int partitions = 1 << partition_bits;
  const int shift = 64 - partition_bits;
  if ((int)dst->size() < partitions) dst->resize(partitions);
  const hmj::detail::KeyColumns cols =
      hmj::detail::flatten<Key, Hash>(begin, end, hmj::detail::get0_of());
  hmj::detail::PinnedArray<uint64_t> hash;
  hmj::detail::PinnedArray<uint32_t> rowid;
  hmj::detail::hash_sort<Key, Hash>(cols, &hash, &rowid);
  // partition sizes from the device counting kernel (reference: leader's reserve(), :62-73)
  std::vector<uint64_t> counts(partitions, 0);
  if (cols.n)
    hmj::Context::check(hmj_partition_counts_host(hmj::Context::get(), hash.data(), cols.n,
                                                  partition_bits, counts.data()),
                        "hmj_partition_counts_host");
  for (int p = 0; p < partitions; ++p) (*dst)[p].reserve((*dst)[p].size() + counts[p]);
  std::vector<const ItemType*> rows;
  rows.reserve(cols.n);
  for (BidirectionalIterator it = begin; it != end; ++it) rows.push_back(&*it);
  for (std::size_t i = 0; i < cols.n; ++i) (*dst)[hash[i] >> shift].push_back(*rows[rowid[i]]);
}

template <typename BidirectionalIterator,
          typename ItemType = typename BidirectionalIterator::value_type,
          typename Key = typename std::tuple_element<0, ItemType>::type,
          typename Value = typename std::tuple_element<1, ItemType>::type,
          typename Hash = std::hash<Key>>
void partitioned_hash_table(BidirectionalIterator begin, BidirectionalIterator end,
                            std::vector<std::unordered_map<Key, Value>>* tables, int /*num_threads*/,
                            int partition_bits) {
  const int partitions = 1 << partition_bits;
  const int shift = 64 - partition_bits;
  if ((int)tables->size() < partitions) tables->resize(partitions);
  const hmj::detail::KeyColumns cols =
      hmj::detail::flatten<Key, Hash>(begin, end, hmj::detail::get0_of());
  hmj::detail::PinnedArray<uint64_t> hash;
  hmj::detail::PinnedArray<uint32_t> rowid;
  hmj::detail::hash_sort<Key, Hash>(cols, &hash, &rowid);
  std::vector<uint64_t> counts(partitions, 0);
  if (cols.n)
    hmj::Context::check(hmj_partition_counts_host(hmj::Context::get(), hash.data(), cols.n,
                                                  partition_bits, counts.data()),
                        "hmj_partition_counts_host");
  for (int p = 0; p < partitions; ++p) (*tables)[p].reserve(counts[p]);
  std::vector<const ItemType*> rows;
  rows.reserve(cols.n);
  for (BidirectionalIterator it = begin; it != end; ++it) rows.push_back(&*it);
  for (std::size_t i = 0; i < cols.n; ++i) {
    const ItemType& item = *rows[rowid[i]];
    (*tables)[hash[i] >> shift].insert(std::make_pair(std::get<0>(item), std::get<1>(item)));
  }
}

}  // namespace radix_hash

#endif  // PARTITIONED_HASH_H
