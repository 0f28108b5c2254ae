// CountMin primitives shared by eval.cu (stand-alone sketch) and sgd_localize.cu (the tail filter
// inside the minibatch localisation).  Hash and probe sequence are the host sketch's
// (csrc/host/util/sketch.h; ref src/util/sketch.h:20-33, countmin.h:21-41).
#pragma once
#include "common.cuh"

namespace psb {

__device__ __forceinline__ uint32_t sketch_hash(uint64_t key) {  // == PS::Sketch::hash on the host
  const uint32_t m = 0xc6a4a793u;
  uint32_t h = 0xbc9f1d34u ^ (8u * m);
  h += (uint32_t)key; h *= m; h ^= h >> 16;
  h += (uint32_t)(key >> 32); h *= m; h ^= h >> 16;
  return h;
}
// saturating add of `count` to byte `pos` of a uint8 table, via CAS on the containing word
__device__ __forceinline__ void sat_add_u8(uint8_t* table, uint32_t pos, uint32_t count, uint32_t vmax) {
  uint32_t* word = reinterpret_cast<uint32_t*>(table) + (pos >> 2);
  const int shift = (pos & 3) * 8;
  uint32_t old = *word, assumed;
  do {
    assumed = old;
    const uint32_t cur = (assumed >> shift) & 0xffu;
    const uint32_t nv = count > vmax - cur ? vmax : cur + count;
    if (nv == cur) return;
    old = atomicCAS(word, assumed, (assumed & ~(0xffu << shift)) | (nv << shift));
  } while (old != assumed);
}

}  // namespace psb
