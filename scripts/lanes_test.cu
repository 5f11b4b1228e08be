// Two host threads, a context each, host-buffer joins back to back (no Python in the picture).
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <thread>
#include <vector>
#include "../include/hmj.h"
static double now() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static inline uint64_t mix(uint64_t x) { x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31); }
int main(int argc, char** argv) {
  const size_t n = size_t(1) << 28;
  const int lanes = argc > 1 ? atoi(argv[1]) : 2;
  uint64_t *r, *s;
  cudaMallocHost(&r, n * 8); cudaMallocHost(&s, n * 8);
  for (size_t i = 0; i < n; ++i) { r[i] = mix(i); s[i] = mix((i * 0x9E3779B1ull) & (n - 1)); }
  std::vector<hmj_ctx*> ctx(lanes);
  std::vector<uint32_t*> o_r(lanes), o_s(lanes);
  for (int i = 0; i < lanes; ++i) {
    hmj_create(&ctx[i], 0);
    cudaMallocHost(&o_r[i], n * 4); cudaMallocHost(&o_s[i], n * 4);
    uint64_t m = 0;
    int rc = hmj_join_u64_host(ctx[i], r, n, s, n, o_r[i], o_s[i], n, &m);
    printf("warm-up lane %d rc %d pairs %llu\n", i, rc, (unsigned long long)m);
  }
  uint64_t m = 0;
  double t0 = now();
  hmj_join_u64_host(ctx[0], r, n, s, n, o_r[0], o_s[0], n, &m);
  const double single = now() - t0;
  printf("single call %.1f ms\n", single);
  std::vector<std::thread> th;
  t0 = now();
  for (int i = 0; i < lanes; ++i)
    th.emplace_back([&, i] {
      std::this_thread::sleep_for(std::chrono::microseconds((long)(i * single * 1000 / lanes)));
      for (int k = 0; k < 4; ++k) {
        uint64_t mm = 0;
        const double a = now();
        hmj_join_u64_host(ctx[i], r, n, s, n, o_r[i], o_s[i], n, &mm);
        printf("lane %d call %d: %.1f -> %.1f (%.1f ms)\n", i, k, a - t0, now() - t0, now() - a);
      }
    });
  for (auto& t : th) t.join();
  printf("%d joins: %.1f ms per join\n", 4 * lanes, (now() - t0) / (4 * lanes));
  return 0;
}
