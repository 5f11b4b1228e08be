// join_device.cuh -- device helpers shared by the join kernels (merge_join.cu, bucket_join.cu).
#pragma once
#include <stdint.h>

// Full-key verification of a candidate pair with string keys: the comparison the reference makes
// after two hashes compare equal (/root/reference hashjoin.h:139-141, key == key).
// Keys start at any byte of their chars column, so the two sides are seldom equally aligned:
// each side is read as the aligned 8-byte words that hold its bytes and shifted into place
// (little endian), 8 key bytes per step.  Only words holding at least one byte of the key are
// touched (device allocations are whole multiples of 8 bytes, so such a word is always mapped).
__device__ __forceinline__ bool keys_equal_str(const uint64_t* __restrict__ aoff,
                                               const uint8_t* __restrict__ achars, uint32_t arow,
                                               const uint64_t* __restrict__ boff,
                                               const uint8_t* __restrict__ bchars, uint32_t brow) {
  const uint64_t a0 = aoff[arow], a1 = aoff[arow + 1];
  const uint64_t b0 = boff[brow], b1 = boff[brow + 1];
  const uint64_t len = a1 - a0;
  if (len != b1 - b0) return false;
  if (len == 0) return true;
  const uintptr_t ua = reinterpret_cast<uintptr_t>(achars + a0);
  const uintptr_t ub = reinterpret_cast<uintptr_t>(bchars + b0);
  const uint64_t* wa = reinterpret_cast<const uint64_t*>(ua & ~uintptr_t(7));
  const uint64_t* wb = reinterpret_cast<const uint64_t*>(ub & ~uintptr_t(7));
  const uint32_t sa = (uint32_t)(ua & 7), sb = (uint32_t)(ub & 7);  // bytes to drop from the first word
  uint64_t lo_a = wa[0], lo_b = wb[0];
  for (uint64_t done = 0, k = 1;; done += 8, ++k) {
    const uint64_t left = len - done;              // key bytes from this step on (>= 1)
    const uint32_t take = left < 8 ? (uint32_t)left : 8u;
    const bool more = left > 8;
    // the step's bytes spill into the next aligned word when the key is not 8-aligned
    const bool hi_a = sa + take > 8, hi_b = sb + take > 8;
    const uint64_t na = (hi_a || more) ? wa[k] : 0, nb = (hi_b || more) ? wb[k] : 0;
    uint64_t va = sa ? (lo_a >> (8 * sa)) | (na << (64 - 8 * sa)) : lo_a;
    uint64_t vb = sb ? (lo_b >> (8 * sb)) | (nb << (64 - 8 * sb)) : lo_b;
    if (take < 8) {
      const uint64_t mask = (uint64_t(1) << (8 * take)) - 1;
      va &= mask;
      vb &= mask;
    }
    if (va != vb) return false;
    if (!more) return true;
    lo_a = na;
    lo_b = nb;
  }
}

// ---- key verification with the loads of several keys in flight ------------------------------
// keys_equal_str walks one pair of keys word by word: offsets, then the first words, then the next
// ones -- two to three dependent memory round trips per candidate, one candidate after the other.
// The merge join's write phase was bound by exactly that latency on string keys.  The helpers
// below split the check into its load stages so that a caller can issue the offset loads of all
// the keys it is about to compare, then all their bytes, and compare at the end.  Keys of up to 32
// bytes are held in registers as the (at most five) aligned words that contain them; longer keys
// take keys_equal_str.
struct hmj_key_span {
  uint64_t start;
  uint64_t len;
};
__device__ __forceinline__ hmj_key_span key_span_of(const uint64_t* __restrict__ off, uint32_t row) {
  hmj_key_span s;
  s.start = off[row];
  s.len = off[row + 1] - s.start;
  return s;
}
struct hmj_key_words {
  uint64_t w[5];
  uint32_t sh;  // bytes to drop from w[0]
};
// requires 1 <= span.len <= 32; touches only words holding at least one byte of the key
__device__ __forceinline__ void key_words_load(const uint8_t* __restrict__ chars, hmj_key_span span,
                                               hmj_key_words* k) {
  const uintptr_t u = reinterpret_cast<uintptr_t>(chars + span.start);
  const uint64_t* wp = reinterpret_cast<const uint64_t*>(u & ~uintptr_t(7));
  k->sh = (uint32_t)(u & 7);
  const uint32_t nwords = (k->sh + (uint32_t)span.len + 7u) >> 3;
#pragma unroll
  for (int i = 0; i < 5; ++i) k->w[i] = (uint32_t)i < nwords ? wp[i] : 0ull;
}
// the key's j-th 8-byte group, little endian (bytes past the key are whatever follows it)
__device__ __forceinline__ uint64_t key_group(const hmj_key_words& k, int j) {
  return k.sh ? (k.w[j] >> (8 * k.sh)) | (k.w[j + 1] << (64 - 8 * k.sh)) : k.w[j];
}
// both keys `len` bytes long (1..32), loaded with key_words_load
__device__ __forceinline__ bool key_words_equal(const hmj_key_words& a, const hmj_key_words& b, uint32_t len) {
  bool eq = true;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if ((uint32_t)(8 * j) < len) {
      uint64_t va = key_group(a, j), vb = key_group(b, j);
      const uint32_t left = len - 8 * j;
      if (left < 8) {
        const uint64_t m = (uint64_t(1) << (8 * left)) - 1;
        va &= m;
        vb &= m;
      }
      eq &= va == vb;
    }
  }
  return eq;
}
