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

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <functional>
#include <iterator>
#include <stdexcept>
#include <new>
#include <string>
#include <thread>
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

// Page-locking memory costs far more than malloc (the driver maps every page), so blocks that a
// call is done with go to a per-thread pool and the next call of similar size takes them back.
class PinnedPool {
 public:
  static void* get(std::size_t bytes, std::size_t* got) {
    Context::get();  // no device: fail with the context's error, not with an allocation failure
    std::vector<Block>& f = blocks();
    std::size_t best = f.size();
    for (std::size_t i = 0; i < f.size(); ++i)
      if (f[i].bytes >= bytes && f[i].bytes <= 2 * bytes + 4096 && (best == f.size() || f[i].bytes < f[best].bytes))
        best = i;
    if (best != f.size()) {
      void* p = f[best].p;
      *got = f[best].bytes;
      f.erase(f.begin() + best);
      return p;
    }
    void* q = nullptr;
    if (hmj_host_alloc(&q, bytes ? bytes : 1) != HMJ_OK) {
      trim(0);  // give the pooled blocks back and try once more
      if (hmj_host_alloc(&q, bytes ? bytes : 1) != HMJ_OK) throw std::bad_alloc();
    }
    *got = bytes;
    return q;
  }
  static void put(void* p, std::size_t bytes) {
    blocks().push_back(Block{p, bytes});
    trim(16);
  }

 private:
  struct Block {
    void* p;
    std::size_t bytes;
  };
  struct Holder {
    std::vector<Block> v;
    ~Holder() {
      for (auto& b : v) hmj_host_free(b.p);
    }
  };
  static std::vector<Block>& blocks() {
    static thread_local Holder h;
    return h.v;
  }
  static void trim(std::size_t keep) {
    std::vector<Block>& f = blocks();
    while (f.size() > keep) {
      hmj_host_free(f.front().p);
      f.erase(f.begin());
    }
  }
};

// Page-locked host array (hmj_host_alloc): what the columns are flattened into and what the
// result pairs arrive in, so that the copies to and from the GPU run at the link's rate and
// asynchronously instead of being staged through the driver's bounce buffers.
template <typename T>
class PinnedArray {
 public:
  PinnedArray() : p_(nullptr), n_(0), cap_(0) {}
  explicit PinnedArray(std::size_t n) : p_(nullptr), n_(0), cap_(0) { resize(n); }
  PinnedArray(PinnedArray&& o) : p_(o.p_), n_(o.n_), cap_(o.cap_), bytes_(o.bytes_) { o.p_ = nullptr; o.n_ = o.cap_ = 0; }
  PinnedArray& operator=(PinnedArray&& o) {
    if (this != &o) { release(); p_ = o.p_; n_ = o.n_; cap_ = o.cap_; bytes_ = o.bytes_; o.p_ = nullptr; o.n_ = o.cap_ = 0; }
    return *this;
  }
  PinnedArray(const PinnedArray&) = delete;
  PinnedArray& operator=(const PinnedArray&) = delete;
  ~PinnedArray() { release(); }
  // contents are NOT preserved when the array grows (every user fills it afterwards)
  void resize(std::size_t n) {
    if (n > cap_) {
      release();
      std::size_t got = 0;
      p_ = static_cast<T*>(PinnedPool::get((n ? n : 1) * sizeof(T), &got));
      bytes_ = got;
      cap_ = got / sizeof(T);
    }
    n_ = n;
  }
  T* data() { return p_; }
  const T* data() const { return p_; }
  std::size_t size() const { return n_; }
  T& operator[](std::size_t i) { return p_[i]; }
  const T& operator[](std::size_t i) const { return p_[i]; }

 private:
  void release() {
    if (p_) PinnedPool::put(p_, bytes_);
    p_ = nullptr;
    n_ = cap_ = 0;
  }
  T* p_;
  std::size_t n_, cap_, bytes_ = 0;
};

// fn(t, lo, hi) over [0, n) cut into `threads` slices, the caller's thread taking the first one.
// This is where the reference's num_threads argument goes: the GPU does the sort and the join,
// the host threads do what is left on the host (copying, flattening, materialising tuples).
template <typename Fn>
void parallel_slices(std::size_t n, int threads, Fn fn) {
  if (threads < 1) threads = 1;
  if ((std::size_t)threads > n / 4096 + 1) threads = (int)(n / 4096 + 1);
  if (threads == 1) {
    fn(0, (std::size_t)0, n);
    return;
  }
  std::vector<std::thread> pool;
  const std::size_t per = (n + threads - 1) / threads;
  for (int t = 1; t < threads; ++t) {
    const std::size_t lo = std::min(n, per * t), hi = std::min(n, per * (t + 1));
    pool.emplace_back([=]() { fn(t, lo, hi); });
  }
  fn(0, (std::size_t)0, std::min(n, per));
  for (auto& th : pool) th.join();
}

// A fixed-size array whose elements are constructed and destroyed by several host threads at once
// (std::vector::resize value-initialises every element on one thread first: 0.1 s for 16M tuples).
template <typename T>
class ParallelArray {
 public:
  ParallelArray() : p_(nullptr), n_(0), threads_(1) {}
  ParallelArray(const ParallelArray&) = delete;
  ParallelArray& operator=(const ParallelArray&) = delete;
  ~ParallelArray() { clear(); }
  // make(i) -> T, called once per element from `threads` threads
  template <typename Make>
  void build(std::size_t n, int threads, Make make) {
    clear();
    p_ = static_cast<T*>(::operator new(sizeof(T) * (n ? n : 1)));
    n_ = n;
    threads_ = threads;
    parallel_slices(n, threads, [&](int, std::size_t lo, std::size_t hi) {
      for (std::size_t i = lo; i < hi; ++i) new (p_ + i) T(make(i));
    });
  }
  void clear() {
    if (!p_) return;
    if (!std::is_trivially_destructible<T>::value) {
      T* p = p_;
      parallel_slices(n_, threads_, [=](int, std::size_t lo, std::size_t hi) {
        for (std::size_t i = lo; i < hi; ++i) p[i].~T();
      });
    }
    ::operator delete(p_);
    p_ = nullptr;
    n_ = 0;
  }
  T* begin() { return p_; }
  T* end() { return p_ + n_; }
  const T* begin() const { return p_; }
  const T* end() const { return p_ + n_; }
  std::size_t size() const { return n_; }
  T& operator[](std::size_t i) { return p_[i]; }
  const T& operator[](std::size_t i) const { return p_[i]; }

 private:
  T* p_;
  std::size_t n_;
  int threads_;
};

// seconds spent in the host-side stages of the last facade call on this thread (facade_bench)
struct Stages {
  double copy_s, flatten_s, device_s, rows_s;
};
inline Stages& stages() {
  static thread_local Stages s = {0, 0, 0, 0};
  return s;
}
inline double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// Columnar copy of the keys of a range of pair/tuple-like items (key = get<0>).
struct KeyColumns {
  PinnedArray<uint64_t> u64;      // kU64: keys as std::hash sees them; kHostHash: Hash{}(key)
  PinnedArray<uint64_t> offsets;  // kString: n+1 offsets
  PinnedArray<uint8_t> chars;     // kString: key bytes (+8 bytes of padding)
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
inline std::size_t key_bytes(const Key&) { return 0; }
inline std::size_t key_bytes(const std::string& s) { return s.size(); }
template <typename Key>
inline void copy_key_bytes(uint8_t*, const Key&) {}
inline void copy_key_bytes(uint8_t* dst, const std::string& s) { std::memcpy(dst, s.data(), s.size()); }

template <typename Iter>
struct is_random_access
    : std::is_base_of<std::random_access_iterator_tag,
                      typename std::iterator_traits<Iter>::iterator_category> {};

// rows [lo, hi) of the range: random access jumps there, anything else walks
template <typename Iter>
inline Iter advance_to(Iter begin, std::size_t lo) {
  std::advance(begin, (typename std::iterator_traits<Iter>::difference_type)lo);
  return begin;
}

// KeyOf: functor returning a const Key& from *iterator.  `threads` host threads share the walk
// when the iterator is random access (a bidirectional range is walked by one thread).
template <typename Key, typename Hash, typename Iter, typename KeyOf>
KeyColumns flatten(Iter begin, Iter end, KeyOf key_of, int threads = 1) {
  const double t0 = now_s();
  KeyColumns c;
  const KeyPath path = key_path<Key, Hash>::value;
  c.n = static_cast<size_t>(std::distance(begin, end));
  if (!is_random_access<Iter>::value) threads = 1;
  if (path == kString) {
    // lengths -> offsets -> bytes
    c.offsets.resize(c.n + 1);
    uint64_t* off = c.offsets.data();
    off[0] = 0;
    parallel_slices(c.n, threads, [&](int, std::size_t lo, std::size_t hi) {
      Iter it = advance_to(begin, lo);
      for (std::size_t i = lo; i < hi; ++i, ++it) off[i + 1] = key_bytes(key_of(*it));
    });
    for (std::size_t i = 0; i < c.n; ++i) off[i + 1] += off[i];
    c.chars.resize((std::size_t)off[c.n] + 8);
    uint8_t* ch = c.chars.data();
    parallel_slices(c.n, threads, [&](int, std::size_t lo, std::size_t hi) {
      Iter it = advance_to(begin, lo);
      for (std::size_t i = lo; i < hi; ++i, ++it) copy_key_bytes(ch + off[i], key_of(*it));
    });
    std::memset(ch + off[c.n], 0, 8);
  } else {
    c.u64.resize(c.n);
    uint64_t* out = c.u64.data();
    parallel_slices(c.n, threads, [&](int, std::size_t lo, std::size_t hi) {
      Hash h;
      Iter it = advance_to(begin, lo);
      for (std::size_t i = lo; i < hi; ++i, ++it)
        out[i] = path == kU64 ? as_u64<Key>(key_of(*it)) : static_cast<uint64_t>(h(key_of(*it)));
    });
  }
  stages().flatten_s += now_s() - t0;
  return c;
}

// (hash,row-id) pairs of a flattened key column, sorted by hash on the GPU (stable)
template <typename Key, typename Hash>
void hash_sort(const KeyColumns& c, PinnedArray<uint64_t>* hash, PinnedArray<uint32_t>* rowid) {
  hash->resize(c.n);
  rowid->resize(c.n);
  if (c.n == 0) return;
  const double t0 = now_s();
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
  stages().device_s += now_s() - t0;
}

// dst[i] = src[perm[i]] for a random-access range, through a temporary
template <typename RandomIt, typename Perm>
void permute_in_place(RandomIt dst, const Perm& perm) {
  typedef typename std::iterator_traits<RandomIt>::value_type T;
  std::vector<T> tmp;
  tmp.reserve(perm.size());
  for (size_t i = 0; i < perm.size(); ++i) tmp.push_back(std::move(dst[perm[i]]));
  for (size_t i = 0; i < perm.size(); ++i) dst[i] = std::move(tmp[i]);
}

}  // namespace detail
}  // namespace hmj

#endif  // HMJ_FACADE_H
