// A from-scratch codec for the Snappy block format (varint length preamble, literal and
// copy elements).  The reference links Google's libsnappy for SArray::compressTo /
// uncompressFrom, the COMPRESSING filter and the slot cache
// (ref src/util/shared_array_inl.h:225-248, src/filter/compressing.h:8-37,
//  src/data/slot_reader.cc:132-143); libsnappy is not on the image, so this implements the
// public format directly.  Streams produced here decode with any Snappy decoder and vice versa.
#pragma once
#include "util/common.h"

namespace PS {
namespace snappy {

inline size_t MaxCompressedLength(size_t n) { return 32 + n + n / 6; }

namespace detail {
inline char* put_varint32(char* p, uint32 v) {
  while (v >= 0x80) { *p++ = (char)(v | 0x80); v >>= 7; }
  *p++ = (char)v;
  return p;
}
inline const char* get_varint32(const char* p, const char* end, uint32* v) {
  uint32 r = 0;
  for (int shift = 0; shift <= 28 && p < end; shift += 7) {
    uint32 b = (uint8)*p++;
    r |= (b & 0x7f) << shift;
    if (!(b & 0x80)) { *v = r; return p; }
  }
  return nullptr;
}
inline char* emit_literal(char* op, const char* lit, size_t len) {
  size_t n = len - 1;
  if (n < 60) {
    *op++ = (char)(n << 2);
  } else {
    char* base = op++;
    int count = 0;
    while (n > 0) { *op++ = (char)(n & 0xff); n >>= 8; ++count; }
    *base = (char)((59 + count) << 2);
  }
  memcpy(op, lit, len);
  return op + len;
}
inline char* emit_copy_upto64(char* op, size_t offset, size_t len) {
  if (len < 12 && offset < 2048) {
    *op++ = (char)(1 | ((len - 4) << 2) | ((offset >> 8) << 5));
    *op++ = (char)(offset & 0xff);
  } else {
    *op++ = (char)(2 | ((len - 1) << 2));
    *op++ = (char)(offset & 0xff);
    *op++ = (char)(offset >> 8);
  }
  return op;
}
inline char* emit_copy(char* op, size_t offset, size_t len) {
  while (len >= 68) { op = emit_copy_upto64(op, offset, 64); len -= 64; }
  if (len > 64) { op = emit_copy_upto64(op, offset, 60); len -= 60; }
  return emit_copy_upto64(op, offset, len);
}
inline uint32 load32(const char* p) { uint32 v; memcpy(&v, p, 4); return v; }
inline uint32 hash32(uint32 v, int shift) { return (v * 0x1e35a7bdu) >> shift; }
}  // namespace detail

// dst must hold MaxCompressedLength(n) bytes
inline void RawCompress(const char* src, size_t n, char* dst, size_t* out_len) {
  using namespace detail;
  char* op = put_varint32(dst, (uint32)n);
  const size_t kBlock = 1 << 16;
  const int kHashBits = 14;
  std::vector<uint16> table(1 << kHashBits);
  for (size_t bstart = 0; bstart < n; bstart += kBlock) {
    const char* base = src + bstart;
    const size_t blen = std::min(kBlock, n - bstart);
    std::fill(table.begin(), table.end(), 0);
    const char* ip = base;
    const char* lit = base;
    const char* end = base + blen;
    if (blen >= 15) {
      const char* limit = end - 15;
      ++ip;
      while (ip < limit) {
        uint32 h = hash32(load32(ip), 32 - kHashBits);
        const char* cand = base + table[h];
        table[h] = (uint16)(ip - base);
        if (cand < ip && load32(cand) == load32(ip)) {
          if (ip > lit) op = emit_literal(op, lit, ip - lit);
          size_t m = 4;
          while (ip + m < end && cand[m] == ip[m]) ++m;
          op = emit_copy(op, ip - cand, m);
          ip += m;
          lit = ip;
          if (ip < limit) table[hash32(load32(ip - 1), 32 - kHashBits)] = (uint16)(ip - 1 - base);
        } else {
          ++ip;
        }
      }
    }
    if (end > lit) op = emit_literal(op, lit, end - lit);
  }
  *out_len = op - dst;
}

inline bool GetUncompressedLength(const char* src, size_t n, size_t* result) {
  uint32 v;
  if (!detail::get_varint32(src, src + n, &v)) return false;
  *result = v;
  return true;
}

// dst must hold GetUncompressedLength() bytes; returns false on a corrupt stream
inline bool RawUncompress(const char* src, size_t n, char* dst) {
  const char* end = src + n;
  uint32 ulen;
  const char* ip = detail::get_varint32(src, end, &ulen);
  if (!ip) return false;
  char* op = dst;
  char* oend = dst + ulen;
  while (ip < end) {
    const uint8 tag = (uint8)*ip++;
    size_t len, offset;
    switch (tag & 3) {
      case 0: {
        len = (tag >> 2) + 1;
        if (len > 60) {
          int nb = (int)len - 60;
          if (ip + nb > end) return false;
          len = 0;
          for (int i = 0; i < nb; ++i) len |= (size_t)(uint8)ip[i] << (8 * i);
          len += 1;
          ip += nb;
        }
        if (ip + len > end || op + len > oend) return false;
        memcpy(op, ip, len);
        ip += len;
        op += len;
        continue;
      }
      case 1:
        if (ip >= end) return false;
        len = ((tag >> 2) & 7) + 4;
        offset = ((size_t)(tag >> 5) << 8) | (uint8)*ip++;
        break;
      case 2:
        if (ip + 2 > end) return false;
        len = (tag >> 2) + 1;
        offset = (uint8)ip[0] | ((size_t)(uint8)ip[1] << 8);
        ip += 2;
        break;
      default:
        if (ip + 4 > end) return false;
        len = (tag >> 2) + 1;
        offset = detail::load32(ip);
        ip += 4;
        break;
    }
    if (offset == 0 || (size_t)(op - dst) < offset || op + len > oend) return false;
    const char* from = op - offset;
    for (size_t i = 0; i < len; ++i) op[i] = from[i];  // may overlap: byte-wise on purpose
    op += len;
  }
  return op == oend;
}

}  // namespace snappy
}  // namespace PS
