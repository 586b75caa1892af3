// Seeded bijection of [0, n): 4-round balanced Feistel network keyed by Philox-2x32-10,
// restricted to [0, n) by cycle walking.  Shared by K7 (csrc/shuffle.cu, the shuffle
// permutation) and K1 (csrc/exchange.cu, distinct sample rows per block).
#pragma once
#include <stdint.h>

namespace xchg_feistel {

__host__ __device__ __forceinline__ uint32_t philox2x32_10(uint32_t c0, uint32_t c1, uint32_t key) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p = (uint64_t)0xD256D193u * c0;
    const uint32_t hi = (uint32_t)(p >> 32), lo = (uint32_t)p;
    c0 = hi ^ key ^ c1;
    c1 = lo;
    key += 0x9E3779B9u;
  }
  return c0;
}

struct Feistel {
  uint32_t key, tweak;  // seed low / high words
  uint32_t half_bits;   // bits per half
  uint64_t half_mask;
  uint64_t n;
};

constexpr int FEISTEL_ROUNDS = 4;

__host__ __device__ __forceinline__ uint64_t feistel_once(const Feistel& f, uint64_t x, bool inverse) {
  uint32_t L = (uint32_t)(x >> f.half_bits), R = (uint32_t)(x & f.half_mask);
  if (!inverse) {
#pragma unroll
    for (int r = 0; r < FEISTEL_ROUNDS; ++r) {
      const uint32_t t = L ^ (philox2x32_10(R, f.tweak ^ (0x9E3779B9u * (uint32_t)(r + 1)), f.key) &
                              (uint32_t)f.half_mask);
      L = R;
      R = t;
    }
  } else {
#pragma unroll
    for (int r = FEISTEL_ROUNDS - 1; r >= 0; --r) {
      const uint32_t t = R ^ (philox2x32_10(L, f.tweak ^ (0x9E3779B9u * (uint32_t)(r + 1)), f.key) &
                              (uint32_t)f.half_mask);
      R = L;
      L = t;
    }
  }
  return ((uint64_t)L << f.half_bits) | R;
}

__host__ __device__ __forceinline__ uint64_t feistel_walk(const Feistel& f, uint64_t x, bool inverse) {
  do {
    x = feistel_once(f, x, inverse);
  } while (x >= f.n);
  return x;
}

__host__ __device__ __forceinline__ void make_feistel(uint64_t seed, uint64_t n_total, Feistel* f) {
  uint32_t bits = 2;
  while (bits < 64 && (1ull << bits) < n_total) ++bits;
  if (bits & 1) ++bits;
  f->key = (uint32_t)seed;
  f->tweak = (uint32_t)(seed >> 32);
  f->half_bits = bits / 2;
  f->half_mask = (1ull << (bits / 2)) - 1;
  f->n = n_total;
}

}  // namespace xchg_feistel
