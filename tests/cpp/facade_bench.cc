// facade_bench.cc -- what a drop-in caller of the C++ facade pays: HashMergeJoin constructor +
// full iteration over std::vector<std::pair<Key, uint64_t>> relations, as hashjoin_bench.cc:109-143
// of dryman/HashMergeJoin times it (BM_HashMergeJoin), with the host-side stages broken out.
//   facade_bench str|u64 <rows> <host threads> [repeats]
// Prints one JSON line.  Needs a B200 (no CPU fallback).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "hashjoin.h"

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// create_strvec's shape (strgen.cc:28-61) over a synthetic word list: q = ceil(sqrt(n)) base
// words, then "w_i-w_j" row-major, shuffled
static std::vector<std::pair<std::string, uint64_t>> make_strings(size_t n, unsigned shuffle_seed) {
  size_t q = 1;
  while (q * q < n) ++q;
  std::mt19937_64 rng(1234);
  std::vector<std::string> words;
  for (size_t i = 0; i < q; ++i) {
    std::string w(3 + rng() % 12, 'a');
    for (auto& c : w) c = (char)('a' + rng() % 26);
    w += std::to_string(i);  // unique
    words.push_back(w);
  }
  std::vector<std::pair<std::string, uint64_t>> v;
  v.reserve(n);
  for (size_t i = 0; i < q && v.size() < n; ++i) v.emplace_back(words[i], i);
  for (size_t i = 0; i < q && v.size() < n; ++i)
    for (size_t j = 0; j < q && v.size() < n; ++j) v.emplace_back(words[i] + "-" + words[j], i + j);
  std::mt19937_64 sh(shuffle_seed);
  std::shuffle(v.begin(), v.end(), sh);
  return v;
}

static std::vector<std::pair<uint64_t, uint64_t>> make_ints(size_t n, unsigned shuffle_seed) {
  std::vector<std::pair<uint64_t, uint64_t>> v(n);
  for (size_t i = 0; i < n; ++i) {
    uint64_t x = i + 0x9E3779B97F4A7C15ull;  // splitmix64 finaliser: distinct keys
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    v[i] = std::make_pair(x ^ (x >> 31), (uint64_t)i);
  }
  std::mt19937_64 sh(shuffle_seed);
  std::shuffle(v.begin(), v.end(), sh);
  return v;
}

template <typename Rel>
static void run(const char* kind, const Rel& r, const Rel& s, unsigned threads, int repeats) {
  typedef typename Rel::const_iterator It;
  double best = 1e30, best_ctor = 0, best_iter = 0, best_dtor = 0;
  hmj::detail::Stages st = {0, 0, 0, 0};
  uint64_t pairs = 0, sum = 0;
  for (int rep = 0; rep < repeats + 1; ++rep) {  // the first call creates the context and grows the workspace
    const double t0 = now_s();
    double t1, t2;
    uint64_t m = 0, acc = 0;
    {
      HashMergeJoin<It, It> join(r.begin(), r.end(), s.begin(), s.end(), threads);
      t1 = now_s();
      for (auto t : join) {  // the consumer of hashjoin_bench.cc:131-133
        acc += *std::get<1>(t) + *std::get<2>(t);
        ++m;
      }
      t2 = now_s();
    }
    const double dt = now_s() - t0;
    if (rep > 0 && dt < best) {
      best = dt;
      best_ctor = t1 - t0;
      best_iter = t2 - t1;
      best_dtor = t0 + dt - t2;
      st = hmj::detail::stages();
    }
    pairs = m;
    sum = acc;
  }
  std::printf("{\"facade\": \"HashMergeJoin ctor + full iteration\", \"keys\": \"%s\", \"rows_per_side\": %zu, "
              "\"host_threads\": %u, \"pairs\": %llu, \"sum\": %llu, \"seconds\": %.6f, \"tuples_per_s\": %.0f, "
              "\"constructor_s\": %.6f, \"iteration_s\": %.6f, \"destructor_s\": %.6f, \"inputs\": \"%s\", "
              "\"stages_s\": {\"copy_relations\": %.6f, \"flatten_to_pinned_columns\": %.6f, "
              "\"device_call_with_copies\": %.6f, \"materialise_tuples\": %.6f}}\n",
              kind, r.size(), threads, (unsigned long long)pairs, (unsigned long long)sum, best,
              2.0 * r.size() / best, best_ctor, best_iter, best_dtor,
#ifdef HMJ_FACADE_BORROW_INPUTS
              "borrowed",
#else
              "copied (reference semantics)",
#endif
              st.copy_s, st.flatten_s, st.device_s, st.rows_s);
}

int main(int argc, char** argv) {
  const char* kind = argc > 1 ? argv[1] : "str";
  const size_t n = argc > 2 ? strtoull(argv[2], nullptr, 10) : 1000000;
  const unsigned threads = argc > 3 ? (unsigned)atoi(argv[3]) : 1;
  const int repeats = argc > 4 ? atoi(argv[4]) : 3;
  if (!strcmp(kind, "str")) {
    auto r = make_strings(n, 1), s = make_strings(n, 2);
    run("strgen strings", r, s, threads, repeats);
  } else {
    auto r = make_ints(n, 1), s = make_ints(n, 2);
    run("u64", r, s, threads, repeats);
  }
  return 0;
}
