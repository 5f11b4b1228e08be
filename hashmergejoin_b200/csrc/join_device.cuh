// join_device.cuh -- device helpers shared by the join kernels (merge_join.cu, bucket_join.cu).
#pragma once
#include <stdint.h>

// Full-key verification of a candidate pair with string keys: the comparison the reference makes
// after two hashes compare equal (/root/reference hashjoin.h:139-141, key == key).
__device__ __forceinline__ bool keys_equal_str(const uint64_t* __restrict__ aoff,
                                               const uint8_t* __restrict__ achars, uint32_t arow,
                                               const uint64_t* __restrict__ boff,
                                               const uint8_t* __restrict__ bchars, uint32_t brow) {
  const uint64_t a0 = aoff[arow], a1 = aoff[arow + 1];
  const uint64_t b0 = boff[brow], b1 = boff[brow + 1];
  const uint64_t len = a1 - a0;
  if (len != b1 - b0) return false;
  const uint8_t* pa = achars + a0;
  const uint8_t* pb = bchars + b0;
  uint64_t i = 0;
  // 8 bytes at a time when both sides are equally aligned, bytes otherwise
  if (((reinterpret_cast<uintptr_t>(pa) ^ reinterpret_cast<uintptr_t>(pb)) & 7) == 0) {
    while (i < len && (reinterpret_cast<uintptr_t>(pa + i) & 7)) {
      if (pa[i] != pb[i]) return false;
      ++i;
    }
    for (; i + 8 <= len; i += 8)
      if (*reinterpret_cast<const uint64_t*>(pa + i) != *reinterpret_cast<const uint64_t*>(pb + i))
        return false;
  }
  for (; i < len; ++i)
    if (pa[i] != pb[i]) return false;
  return true;
}

