// hash_device.cuh -- device restatement of libstdc++'s 64-bit std::_Hash_bytes
// (GCC 13.3, libstdc++-v3/libsupc++/hash_bytes.cc; the function std::hash<std::string> calls
// with seed 0xc70f6907, see /usr/include/c++/13/bits/functional_hash.h).  It is the hash the
// reference uses by default (radix_hash.h:353 `typename Hash = std::hash<Key>`).
#pragma once
#include <stdint.h>

namespace hmj_dev {

constexpr uint64_t kStdHashSeed = 0xc70f6907ULL;
constexpr uint64_t kMul = 0xc6a4a7935bd1e995ULL;  // (0xc6a4a793 << 32) + 0x5bd1e995

__device__ __forceinline__ uint64_t shift_mix(uint64_t v) { return v ^ (v >> 47); }

// Hashes `len` bytes that start at byte offset `byte_start` of an 8-byte-aligned word array.
// Words are read aligned; a row that starts mid-word is reassembled with a funnel shift and a
// one-word carry, so every byte is loaded once.  GUARD=true never touches a word past the one
// holding the last byte (global memory); GUARD=false may read one word further (padded smem).
template <bool GUARD, typename LenT>
__device__ __forceinline__ uint64_t hash_bytes_words(const uint64_t* __restrict__ words,
                                                     uint64_t byte_start, LenT len,
                                                     int /*unused*/) {
  uint64_t h = kStdHashSeed ^ ((uint64_t)len * kMul);
  if (len != 0) {
    const uint32_t sh = (uint32_t)(byte_start & 7) * 8;
    uint64_t widx = byte_start >> 3;
    const uint64_t last_word = (byte_start + (uint64_t)len - 1) >> 3;
    uint64_t cur = words[widx];
    const LenT nwords = len >> 3;
    for (LenT k = 0; k < nwords; ++k) {
      ++widx;
      uint64_t nxt = 0;
      if (!GUARD || widx <= last_word) nxt = words[widx];
      const uint64_t w = sh ? ((cur >> sh) | (nxt << (64 - sh))) : cur;
      cur = nxt;
      const uint64_t d = shift_mix(w * kMul) * kMul;
      h ^= d;
      h *= kMul;
    }
    const uint32_t tail = (uint32_t)(len & 7);
    if (tail) {
      uint64_t w = cur >> sh;
      if (sh && sh + tail * 8 > 64) {
        // tail straddles into the next word
        uint64_t nxt = 0;
        if (!GUARD || widx + 1 <= last_word) nxt = words[widx + 1];
        w |= nxt << (64 - sh);
      }
      w &= (~0ULL) >> (64 - tail * 8);
      h ^= w;
      h *= kMul;
    }
  }
  h = shift_mix(h) * kMul;
  h = shift_mix(h);
  return h;
}

}  // namespace hmj_dev
