// facade_test.cc -- the reference's unit tests, restated against the facade headers
// (hashmergejoin_b200/facade/*.h -> libhmj.so -> CUDA).  Cases follow radix_hash_test.cc:40-262,
// radix_sort_test.cc:27-68, partitioned_hash_test.cc:14-46 and thread_barrier_test.cc:23-55 of
// dryman/HashMergeJoin; the reference has no test for hashjoin.h, so the join cases compare with
// a brute-force equi-join.  Needs a B200: there is no CPU fallback.  Exit code = failures.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <map>
#include <random>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

#include "hashjoin.h"
#include "partitioned_hash.h"
#include "radix_hash.h"
#include "radix_sort.h"
#include "thread_barrier.h"

static int g_fail = 0, g_checks = 0;
#define CHECK(cond)                                                          \
  do {                                                                       \
    ++g_checks;                                                              \
    if (!(cond)) {                                                           \
      ++g_fail;                                                              \
      std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond);            \
    }                                                                        \
  } while (0)

struct identity_hash {
  std::size_t operator()(const int& k) const { return k; }
  std::size_t operator()(const uint64_t& k) const { return k; }
};

typedef std::tuple<std::size_t, int, int> IntTuple;

template <typename It>
static bool hashes_ascending(It b, It e) {
  for (It it = b; it != e; ++it)
    if (it != b && std::get<0>(*(it - 1)) > std::get<0>(*it)) return false;
  return true;
}

// radix_non_inplace_par ladder: 5 reversed (bits 3, 1..2), 12345 reversed (bits 1, 14), 2^18 auto
static void test_non_inplace() {
  for (int bits : {3, 1, 2}) {
    std::vector<std::pair<int, int>> src;
    for (int i = 5; i > 0; i--) src.push_back(std::make_pair(i, i));
    std::vector<IntTuple> dst(5);
    radix_hash::radix_non_inplace_par<int, int, identity_hash>(src.begin(), src.end(), dst.begin(), 1, bits);
    for (int i = 0; i < 5; i++) CHECK(std::get<0>(dst[i]) == (std::size_t)(i + 1));
    for (int i = 0; i < 5; i++) CHECK(std::get<1>(dst[i]) == i + 1 && std::get<2>(dst[i]) == i + 1);
  }
  for (int bits : {1, 14}) {
    const int n = 12345;
    std::vector<std::pair<int, int>> src;
    for (int i = n; i > 0; i--) src.push_back(std::make_pair(i, i));
    std::vector<IntTuple> dst(n);
    radix_hash::radix_non_inplace_par<int, int, identity_hash>(
        src.begin(), src.end(), dst.begin(), std::thread::hardware_concurrency(), bits);
    bool ok = true;
    for (int i = 0; i < n; i++) ok &= std::get<0>(dst[i]) == (std::size_t)(i + 1);
    CHECK(ok);
  }
  {
    const int n = 1 << 18;
    std::vector<std::pair<int, int>> src;
    for (int i = n; i > 0; i--) src.push_back(std::make_pair(i, i));
    std::vector<IntTuple> dst(n);
    radix_hash::radix_non_inplace_par<int, int, identity_hash>(src.begin(), src.end(), dst.begin(),
                                                              std::thread::hardware_concurrency());
    bool ok = true;
    for (int i = 0; i < n; i++) ok &= std::get<0>(dst[i]) == (std::size_t)(i + 1);
    CHECK(ok);
  }
  {  // default Hash = std::hash<std::string>: hash column equals the host's std::hash, ascending
    std::vector<std::pair<std::string, uint64_t>> src;
    for (uint64_t i = 0; i < 5000; i++) src.push_back(std::make_pair("key-" + std::to_string(i * 7919), i));
    std::vector<std::tuple<std::size_t, std::string, uint64_t>> dst(src.size());
    radix_hash::radix_non_inplace_par<std::string, uint64_t>(src.begin(), src.end(), dst.begin(), 4);
    CHECK(hashes_ascending(dst.begin(), dst.end()));
    bool ok = true;
    for (auto& t : dst)
      ok &= std::get<0>(t) == std::hash<std::string>()(std::get<1>(t)) &&
            std::get<1>(t) == "key-" + std::to_string(std::get<2>(t) * 7919);
    CHECK(ok);
  }
}

static void test_inplace() {
  for (int bits : {3, 1, 2}) {
    std::vector<IntTuple> dst;
    for (int i = 5; i > 0; i--) dst.push_back(std::make_tuple((std::size_t)i, i, i));
    radix_hash::radix_inplace_seq<int, int>(dst.begin(), 5, bits);
    for (int i = 0; i < 5; i++) CHECK(std::get<0>(dst[i]) == (std::size_t)(i + 1) && std::get<1>(dst[i]) == i + 1);
  }
  {  // 2^16 random u64 hashes against std::sort (radix_hash_test.cc:195-212)
    std::mt19937_64 rng(7);
    std::vector<std::tuple<std::size_t, uint64_t, uint64_t>> a;
    for (int i = 0; i < (1 << 16); i++) {
      uint64_t r = rng();
      a.push_back(std::make_tuple((std::size_t)r, r, r));
    }
    auto b = a;
    std::sort(b.begin(), b.end());
    radix_hash::radix_inplace_seq<uint64_t, uint64_t>(a.begin(), a.size());
    CHECK(a == b);
  }
  {  // parallel entry points: 2^18 reversed with 4 threads; top-bit values with 8 threads, bits=1
    const int n = 1 << 18;
    std::vector<IntTuple> dst;
    for (int i = n; i > 0; i--) dst.push_back(std::make_tuple((std::size_t)i, i, i));
    radix_hash::radix_inplace_par(dst.begin(), n, 4);
    bool ok = true;
    for (int i = 0; i < n; i++) ok &= std::get<0>(dst[i]) == (std::size_t)(i + 1);
    CHECK(ok);
    std::vector<std::tuple<std::size_t, uint64_t, uint64_t>> t;
    const uint64_t top = 1ULL << 63;
    for (uint64_t i = 0; i < 512; i++) {
      t.push_back(std::make_tuple((std::size_t)(i | top), i, i));
      t.push_back(std::make_tuple((std::size_t)i, i, i));
    }
    radix_hash::radix_inplace_par(t.begin(), t.size(), 8, 1);
    CHECK(hashes_ascending(t.begin(), t.end()));
    CHECK(std::get<0>(t[0]) == 0 && std::get<0>(t[1023]) == (std::size_t)(top | 511));
  }
}

static void test_radix_sort() {  // radix_sort_test.cc:27-68
  std::mt19937_64 rng(11);
  std::vector<std::pair<uint64_t, uint64_t>> a;
  for (int i = 0; i < (1 << 16); i++) {
    uint64_t r = rng();
    a.push_back(std::make_pair(r, r ^ 5));
  }
  auto expect = a;
  std::sort(expect.begin(), expect.end());
  auto in_place = a;
  ::radix_int_inplace<uint64_t, uint64_t>(in_place.begin(), in_place.size(), 4);
  CHECK(in_place == expect);
  std::vector<std::pair<uint64_t, uint64_t>> out(a.size());
  ::radix_int_non_inplace<uint64_t, uint64_t>(a.begin(), a.end(), out.begin(), 4);
  CHECK(out == expect);
}

static void test_partition() {  // partitioned_hash_test.cc:14-46
  typedef std::pair<uint64_t, uint64_t> Item;
  std::vector<Item> src;
  const uint64_t top = 1ULL << 63;
  for (uint64_t i = 0; i < 5; i++) {
    src.push_back(std::make_pair(i, i));
    src.push_back(std::make_pair(i | top, i | top));
    src.push_back(std::make_pair(i | top | 1024, i | top | 1024));
  }
  std::vector<std::vector<Item>> dst(2);
  radix_hash::partition_only<decltype(src.begin()), Item, uint64_t, uint64_t, identity_hash>(
      src.begin(), src.end(), &dst, 2, 1);
  CHECK(dst[0].size() == 5);
  CHECK(dst[1].size() == 10);
  bool ok = true;
  for (auto& it : dst[0]) ok &= (it.first >> 63) == 0;
  for (auto& it : dst[1]) ok &= (it.first >> 63) == 1;
  CHECK(ok);
  std::vector<std::unordered_map<uint64_t, uint64_t>> tables(2);
  radix_hash::partitioned_hash_table<decltype(src.begin()), Item, uint64_t, uint64_t, identity_hash>(
      src.begin(), src.end(), &tables, 1, 1);
  CHECK(tables[0].size() == 5);
  CHECK(tables[1].size() == 10);
  CHECK(tables[1].at(3 | top | 1024) == (3 | top | 1024));
  // default std::hash<std::string>, 10 bits (hashjoin_bench.cc:88-89)
  std::vector<std::pair<std::string, uint64_t>> strs;
  for (uint64_t i = 0; i < 20000; i++) strs.push_back(std::make_pair("w" + std::to_string(i) + "-x", i));
  std::vector<std::vector<std::pair<std::string, uint64_t>>> parts(1024);
  radix_hash::partition_only(strs.begin(), strs.end(), &parts, 4, 10);
  std::size_t total = 0;
  ok = true;
  for (int p = 0; p < 1024; p++) {
    total += parts[p].size();
    for (auto& it : parts[p]) ok &= (std::hash<std::string>()(it.first) >> 54) == (std::size_t)p;
  }
  CHECK(total == strs.size());
  CHECK(ok);
}

static void test_barrier() {  // thread_barrier_test.cc:23-55
  const int n = 20;
  ThreadBarrier barrier(n);
  std::atomic_int leaders(0), followers(0);
  std::vector<std::thread> th;
  for (int i = 0; i < n; i++)
    th.emplace_back([&] { (barrier.wait() ? leaders : followers)++; });
  for (auto& t : th) t.join();
  CHECK(leaders == 1);
  CHECK(followers == n - 1);
}

template <typename K>
static std::vector<std::tuple<K, uint64_t, uint64_t>> brute_join(
    const std::vector<std::pair<K, uint64_t>>& r, const std::vector<std::pair<K, uint64_t>>& s) {
  std::multimap<K, uint64_t> m;
  for (auto& x : s) m.insert(x);
  std::vector<std::tuple<K, uint64_t, uint64_t>> out;
  for (auto& x : r) {
    auto range = m.equal_range(x.first);
    for (auto it = range.first; it != range.second; ++it) out.push_back(std::make_tuple(x.first, x.second, it->second));
  }
  std::sort(out.begin(), out.end());
  return out;
}

static void test_hashjoin() {
  {  // hashjoin_bench.cc:109-143 shape: same string key set on both sides, two orders, sum consumer
    KeyValVec r, s;
    for (uint64_t i = 0; i < 30000; i++) r.push_back(std::make_pair("word" + std::to_string(i) + "-w" + std::to_string(i % 173), i));
    s = r;
    std::mt19937 g(3);
    std::shuffle(s.begin(), s.end(), g);
    for (auto& x : s) x.second += 1000000;
    HashMergeJoin<KeyValVec::iterator, KeyValVec::iterator> hmj(r.begin(), r.end(), s.begin(), s.end(), 4);
    std::vector<std::tuple<std::string, uint64_t, uint64_t>> got;
    uint64_t sum = 0;
    for (auto t : hmj) {
      sum += *std::get<1>(t) + *std::get<2>(t);
      got.push_back(std::make_tuple(*std::get<0>(t), *std::get<1>(t), *std::get<2>(t)));
    }
    std::sort(got.begin(), got.end());
    CHECK(got == brute_join(r, s));
    uint64_t expect = 0;
    for (auto& x : r) expect += 2 * x.second + 1000000;
    CHECK(sum == expect);
    hmj.clear();
    CHECK(hmj.begin() == hmj.end());
  }
  {  // u64 keys with duplicates on both sides: full cross product of every run
    std::vector<std::pair<uint64_t, uint64_t>> r, s;
    std::mt19937_64 rng(5);
    for (uint64_t i = 0; i < 20000; i++) r.push_back(std::make_pair(rng() % 3000, i));
    for (uint64_t i = 0; i < 15000; i++) s.push_back(std::make_pair(rng() % 3000, i));
    typedef std::vector<std::pair<uint64_t, uint64_t>>::iterator It;
    HashMergeJoin<It, It> hmj(r.begin(), r.end(), s.begin(), s.end());
    std::vector<std::tuple<uint64_t, uint64_t, uint64_t>> got;
    for (auto it = hmj.begin(); it != hmj.end(); ++it)
      got.push_back(std::make_tuple(*std::get<0>(*it), *std::get<1>(*it), *std::get<2>(*it)));
    std::sort(got.begin(), got.end());
    CHECK(got == brute_join(r, s));
  }
  {  // HashMergeJoin2: pre-hashed tuples, sorted in place, then joined (hashjoin_bench.cc:180-223)
    HashKeyValVec r, s;
    KeyValVec rk, sk;
    for (uint64_t i = 0; i < 8000; i++) {
      std::string k = "k" + std::to_string(i * 31 % 5000);
      r.push_back(std::make_tuple(std::hash<std::string>()(k), k, i));
      rk.push_back(std::make_pair(k, i));
    }
    for (uint64_t i = 0; i < 6000; i++) {
      std::string k = "k" + std::to_string(i * 17 % 5000);
      s.push_back(std::make_tuple(std::hash<std::string>()(k), k, i + 77));
      sk.push_back(std::make_pair(k, i + 77));
    }
    HashMergeJoin2<HashKeyValVec::iterator, HashKeyValVec::iterator> hmj(r.begin(), r.end(), s.begin(), s.end(), 2);
    CHECK(hashes_ascending(r.begin(), r.end()));
    CHECK(hashes_ascending(s.begin(), s.end()));
    std::vector<std::tuple<std::string, uint64_t, uint64_t>> got;
    for (auto t : hmj) got.push_back(std::make_tuple(*std::get<0>(t), *std::get<1>(t), *std::get<2>(t)));
    std::sort(got.begin(), got.end());
    CHECK(got == brute_join(rk, sk));
  }
  {  // empty sides
    KeyValVec r, s;
    r.push_back(std::make_pair("a", 1));
    HashMergeJoin<KeyValVec::iterator, KeyValVec::iterator> a(r.begin(), r.end(), s.begin(), s.end());
    CHECK(a.begin() == a.end());
    HashMergeJoin<KeyValVec::iterator, KeyValVec::iterator> b(s.begin(), s.end(), r.begin(), r.end());
    CHECK(b.begin() == b.end());
  }
}

int main() {
  CHECK(radix_hash::optimal_partition(1000000) == 10);  // SURVEY.md 8a A2 [probe]: 1M -> 10
  CHECK(radix_hash::optimal_partition(64u << 20) == 13);
  CHECK(radix_hash::optimal_partition(10) == 6);
  test_barrier();
  try {
    test_non_inplace();
    test_inplace();
    test_radix_sort();
    test_partition();
    test_hashjoin();
  } catch (const std::exception& e) {
    std::printf("FAIL exception: %s\n", e.what());
    ++g_fail;
  }
  std::printf("facade_test: %d checks, %d failures\n", g_checks, g_fail);
  return g_fail ? 1 : 0;
}
