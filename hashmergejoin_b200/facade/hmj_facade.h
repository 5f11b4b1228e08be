// hmj_facade.h -- shared plumbing of the C++ facade headers (radix_hash.h, hashjoin.h,
// partitioned_hash.h, radix_sort.h).  The facade keeps the names, template parameters and
// argument order of dryman/HashMergeJoin's header templates and forwards the work to the C ABI
// of libhmj.so (include/hmj.h); nothing in here computes a hash order or a join on the host.
//
// Host-side work that remains, by necessity: walking the caller's iterators once to flatten keys
// into columns (u64 keys, or offsets + chars for std::string), and materialising result tuples
// from the row ids the GPU returns.
#ifndef HMJ_FACADE_H
#define HMJ_FACADE_H 1

#include <cstdint>
#include <cstring>
#include <functional>
#include <iterator>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "hmj.h"

namespace hmj {

// One context (device 0, own stream) per thread, created on first use; destroyed at thread exit.
class Context {
 public:
  static hmj_ctx* get() {
    static thread_local Context c;
    return c.ctx_;
  }
  static void check(int rc, const char* what) {
    if (rc == HMJ_OK) return;
    std::string msg = std::string(what) + ": ";
    msg += hmj_last_error(get());
    throw std::runtime_error(msg);  // no CPU fallback: errors surface to the caller
  }

 private:
  Context() : ctx_(nullptr) {
    int rc = hmj_create(&ctx_, 0);
    if (rc != HMJ_OK)
      throw std::runtime_error("hmj_create failed (no sm_100 device?): code " + std::to_string(rc));
  }
  ~Context() {
    if (ctx_) hmj_destroy(ctx_);
  }
  hmj_ctx* ctx_;
};

namespace detail {

// Which device path a (Key, Hash) combination takes.
//   kU64    : Hash is std::hash of an integral key -> identity, key column = hash column
//   kString : Hash is std::hash<std::string>       -> _Hash_bytes on the device, keys verified
//   kHostHash : any other functor is user code that cannot run on the device; it is evaluated
//               while flattening and the hash column is shipped (sort/partition/join on the GPU)
enum KeyPath { kU64, kString, kHostHash };

template <typename Key, typename Hash>
struct key_path {
  static const KeyPath value =
      (std::is_same<Hash, std::hash<Key>>::value && std::is_integral<Key>::value)
          ? kU64
          : (std::is_same<Hash, std::hash<Key>>::value && std::is_same<Key, std::string>::value)
                ? kString
                : kHostHash;
};

// Columnar copy of the keys of a range of pair/tuple-like items (key = get<0>).
struct KeyColumns {
  std::vector<uint64_t> u64;      // kU64: keys as std::hash sees them; kHostHash: Hash{}(key)
  std::vector<uint64_t> offsets;  // kString: n+1 offsets
  std::vector<uint8_t> chars;     // kString: key bytes (+8 bytes of padding)
  size_t n = 0;
};

template <typename Key>
inline typename std::enable_if<std::is_integral<Key>::value, uint64_t>::type as_u64(const Key& k) {
  return static_cast<uint64_t>(std::hash<Key>()(k));  // libstdc++: static_cast<size_t>(k)
}
template <typename Key>
inline typename std::enable_if<!std::is_integral<Key>::value, uint64_t>::type as_u64(const Key&) {
  return 0;
}
template <typename Key>
inline void append_chars(KeyColumns* c, const Key&) { (void)c; }
inline void append_chars(KeyColumns* c, const std::string& s) {
  c->chars.insert(c->chars.end(), s.begin(), s.end());
  c->offsets.push_back(c->chars.size());
}

// KeyOf: functor returning a const Key& from *iterator
template <typename Key, typename Hash, typename Iter, typename KeyOf>
KeyColumns flatten(Iter begin, Iter end, KeyOf key_of) {
  KeyColumns c;
  const KeyPath path = key_path<Key, Hash>::value;
  c.n = static_cast<size_t>(std::distance(begin, end));
  if (path == kString) {
    c.offsets.reserve(c.n + 1);
    c.offsets.push_back(0);
    for (Iter it = begin; it != end; ++it) append_chars(&c, key_of(*it));
    c.chars.resize(c.chars.size() + 8, 0);
  } else {
    c.u64.reserve(c.n);
    Hash h;
    for (Iter it = begin; it != end; ++it) {
      if (path == kU64) c.u64.push_back(as_u64<Key>(key_of(*it)));
      else c.u64.push_back(static_cast<uint64_t>(h(key_of(*it))));
    }
  }
  return c;
}

// (hash,row-id) pairs of a flattened key column, sorted by hash on the GPU (stable)
template <typename Key, typename Hash>
void hash_sort(const KeyColumns& c, std::vector<uint64_t>* hash, std::vector<uint32_t>* rowid) {
  hash->resize(c.n);
  rowid->resize(c.n);
  if (c.n == 0) return;
  hmj_ctx* ctx = Context::get();
  if (key_path<Key, Hash>::value == kString)
    Context::check(hmj_hash_sort_strings_host(ctx, c.offsets.data(), c.chars.data(), c.n,
                                              hash->data(), rowid->data()), "hmj_hash_sort_strings_host");
  else if (key_path<Key, Hash>::value == kU64)
    Context::check(hmj_hash_sort_u64_host(ctx, c.u64.data(), c.n, hash->data(), rowid->data()),
                   "hmj_hash_sort_u64_host");
  else
    Context::check(hmj_sort_pairs_host(ctx, c.u64.data(), nullptr, c.n, hash->data(), rowid->data()),
                   "hmj_sort_pairs_host");
}

// dst[i] = src[perm[i]] for a random-access range, through a temporary
template <typename RandomIt>
void permute_in_place(RandomIt dst, const std::vector<uint32_t>& perm) {
  typedef typename std::iterator_traits<RandomIt>::value_type T;
  std::vector<T> tmp;
  tmp.reserve(perm.size());
  for (size_t i = 0; i < perm.size(); ++i) tmp.push_back(std::move(dst[perm[i]]));
  for (size_t i = 0; i < perm.size(); ++i) dst[i] = std::move(tmp[i]);
}

}  // namespace detail
}  // namespace hmj

#endif  // HMJ_FACADE_H
