// CRC32C (Castagnoli) and MurmurHash3, written from the algorithm definitions.
// ref src/util/crc32c.{h,cc} (LevelDB table-driven CRC used for key-cache signatures,
// src/filter/key_caching.h:18,74) and src/util/murmurhash3.{h,cc} (feature-id shuffling in
// the TERAFEA parser, src/data/text_parser.cc:129-181).
#pragma once
#include "util/common.h"

namespace PS {
namespace crc32c {

// reflected polynomial 0x82F63B78; 8 slice tables built once
struct Tables {
  uint32 t[8][256];
  Tables() {
    for (uint32 i = 0; i < 256; ++i) {
      uint32 c = i;
      for (int k = 0; k < 8; ++k) c = (c >> 1) ^ (0x82F63B78u & (0u - (c & 1u)));
      t[0][i] = c;
    }
    for (uint32 i = 0; i < 256; ++i)
      for (int s = 1; s < 8; ++s) t[s][i] = (t[s - 1][i] >> 8) ^ t[0][t[s - 1][i] & 0xff];
  }
};
inline const Tables& tables() {
  static const Tables tb;
  return tb;
}

// crc of concat(A, data) given crc(A) = init_crc
inline uint32 Extend(uint32 init_crc, const char* data, size_t n) {
  const Tables& tb = tables();
  const uint8* p = reinterpret_cast<const uint8*>(data);
  uint32 c = init_crc ^ 0xffffffffu;
  while (n && (reinterpret_cast<uintptr_t>(p) & 7)) { c = (c >> 8) ^ tb.t[0][(c ^ *p++) & 0xff]; --n; }
  while (n >= 8) {
    uint64 w;
    memcpy(&w, p, 8);
    w ^= c;
    c = tb.t[7][w & 0xff] ^ tb.t[6][(w >> 8) & 0xff] ^ tb.t[5][(w >> 16) & 0xff] ^
        tb.t[4][(w >> 24) & 0xff] ^ tb.t[3][(w >> 32) & 0xff] ^ tb.t[2][(w >> 40) & 0xff] ^
        tb.t[1][(w >> 48) & 0xff] ^ tb.t[0][(w >> 56) & 0xff];
    p += 8;
    n -= 8;
  }
  while (n--) c = (c >> 8) ^ tb.t[0][(c ^ *p++) & 0xff];
  return c ^ 0xffffffffu;
}
inline uint32 Value(const char* data, size_t n) { return Extend(0, data, n); }

// masked form for storing a crc next to the data it covers (LevelDB convention)
static const uint32 kMaskDelta = 0xa282ead8ul;
inline uint32 Mask(uint32 crc) { return ((crc >> 15) | (crc << 17)) + kMaskDelta; }
inline uint32 Unmask(uint32 masked) {
  uint32 rot = masked - kMaskDelta;
  return (rot >> 17) | (rot << 15);
}
}  // namespace crc32c

// ---- MurmurHash3 (x86_32, x64_128) -------------------------------------------------------------
inline uint32 rotl32(uint32 x, int8 r) { return (x << r) | (x >> (32 - r)); }
inline uint64 rotl64(uint64 x, int8 r) { return (x << r) | (x >> (64 - r)); }
inline uint32 fmix32(uint32 h) {
  h ^= h >> 16; h *= 0x85ebca6b; h ^= h >> 13; h *= 0xc2b2ae35; h ^= h >> 16;
  return h;
}
inline uint64 fmix64(uint64 k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
  return k;
}

inline void MurmurHash3_x86_32(const void* key, int len, uint32 seed, void* out) {
  const uint8* data = static_cast<const uint8*>(key);
  const int nblocks = len / 4;
  uint32 h1 = seed;
  const uint32 c1 = 0xcc9e2d51, c2 = 0x1b873593;
  for (int i = 0; i < nblocks; ++i) {
    uint32 k1;
    memcpy(&k1, data + 4 * i, 4);
    k1 *= c1; k1 = rotl32(k1, 15); k1 *= c2;
    h1 ^= k1; h1 = rotl32(h1, 13); h1 = h1 * 5 + 0xe6546b64;
  }
  const uint8* tail = data + nblocks * 4;
  uint32 k1 = 0;
  switch (len & 3) {
    case 3: k1 ^= tail[2] << 16;  // fallthrough
    case 2: k1 ^= tail[1] << 8;   // fallthrough
    case 1: k1 ^= tail[0]; k1 *= c1; k1 = rotl32(k1, 15); k1 *= c2; h1 ^= k1;
  }
  h1 ^= len;
  *static_cast<uint32*>(out) = fmix32(h1);
}

inline void MurmurHash3_x64_128(const void* key, int len, uint32 seed, void* out) {
  const uint8* data = static_cast<const uint8*>(key);
  const int nblocks = len / 16;
  uint64 h1 = seed, h2 = seed;
  const uint64 c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
  for (int i = 0; i < nblocks; ++i) {
    uint64 k1, k2;
    memcpy(&k1, data + 16 * i, 8);
    memcpy(&k2, data + 16 * i + 8, 8);
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
    h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
    k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
    h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
  }
  const uint8* tail = data + nblocks * 16;
  uint64 k1 = 0, k2 = 0;
  switch (len & 15) {
    case 15: k2 ^= (uint64)tail[14] << 48;  // fallthrough
    case 14: k2 ^= (uint64)tail[13] << 40;  // fallthrough
    case 13: k2 ^= (uint64)tail[12] << 32;  // fallthrough
    case 12: k2 ^= (uint64)tail[11] << 24;  // fallthrough
    case 11: k2 ^= (uint64)tail[10] << 16;  // fallthrough
    case 10: k2 ^= (uint64)tail[9] << 8;    // fallthrough
    case 9: k2 ^= (uint64)tail[8]; k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;  // fallthrough
    case 8: k1 ^= (uint64)tail[7] << 56;    // fallthrough
    case 7: k1 ^= (uint64)tail[6] << 48;    // fallthrough
    case 6: k1 ^= (uint64)tail[5] << 40;    // fallthrough
    case 5: k1 ^= (uint64)tail[4] << 32;    // fallthrough
    case 4: k1 ^= (uint64)tail[3] << 24;    // fallthrough
    case 3: k1 ^= (uint64)tail[2] << 16;    // fallthrough
    case 2: k1 ^= (uint64)tail[1] << 8;     // fallthrough
    case 1: k1 ^= (uint64)tail[0]; k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
  }
  h1 ^= len; h2 ^= len;
  h1 += h2; h2 += h1;
  h1 = fmix64(h1); h2 = fmix64(h2);
  h1 += h2; h2 += h1;
  static_cast<uint64*>(out)[0] = h1;
  static_cast<uint64*>(out)[1] = h2;
}

}  // namespace PS
