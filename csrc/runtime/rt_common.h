// Shared helpers of libtfdl_runtime.so (host-only C++).
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace tfdl_rt {

void set_error(const char* fmt, ...);
const char* last_error();

uint32_t crc32c_extend(uint32_t crc, const void* data, size_t n);
inline uint32_t crc32c_value(const void* data, size_t n) { return crc32c_extend(0, data, n); }
// TensorFlow / LevelDB "masked" crc: rotate right by 15 and add a constant.
inline uint32_t crc32c_mask(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }
inline uint32_t crc32c_unmask(uint32_t m) {
  uint32_t rot = m - 0xa282ead8u;
  return (rot >> 17) | (rot << 15);
}

inline void put_varint64(std::string* s, uint64_t v) {
  while (v >= 128) {
    s->push_back(static_cast<char>(v | 128));
    v >>= 7;
  }
  s->push_back(static_cast<char>(v));
}
inline void put_varint32(std::string* s, uint32_t v) { put_varint64(s, v); }
inline void put_fixed32(std::string* s, uint32_t v) { s->append(reinterpret_cast<const char*>(&v), 4); }
inline void put_fixed64(std::string* s, uint64_t v) { s->append(reinterpret_cast<const char*>(&v), 8); }
inline bool get_varint64(const uint8_t*& p, const uint8_t* end, uint64_t* out) {
  uint64_t r = 0;
  for (int shift = 0; shift <= 63 && p < end; shift += 7) {
    uint64_t b = *p++;
    r |= (b & 127) << shift;
    if (!(b & 128)) {
      *out = r;
      return true;
    }
  }
  return false;
}

// splitmix64 / xoshiro256** - deterministic across platforms (the Python fallback does not need to
// match bit for bit; shuffles only have to be valid permutations with the tf.data buffer semantics).
struct Rng {
  uint64_t s[4];
  explicit Rng(uint64_t seed) {
    for (auto& x : s) {
      seed += 0x9e3779b97f4a7c15ull;
      uint64_t z = seed;
      z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
      z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
      x = z ^ (z >> 31);
    }
  }
  static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
  uint64_t next() {
    uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return r;
  }
  // unbiased integer in [0, n)
  uint64_t below(uint64_t n) {
    uint64_t x, lim = UINT64_MAX - UINT64_MAX % n;
    do { x = next(); } while (x >= lim);
    return x % n;
  }
};

}  // namespace tfdl_rt
