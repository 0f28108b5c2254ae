// Dense feature-block tile on the 5th-generation tensor cores (the one GEMM-shaped op of this
// framework): margins of n examples with d DENSE features against K weight vectors,
//
//     Y[n x K] = X[n x d] * W[K x d]^T          (bf16 operands, fp32 accumulation in TMEM)
//
// with the logistic loss of every (example, model) pair reduced per model in the epilogue, so a
// set of K candidate models (a regularisation path, the K shards of a one-vs-rest problem) is
// scored on a dense slot in one pass over X.  PS-format DENSE slots (ref
// src/data/text_parser.cc:27,200-250) are the only dense matrices the reference knows; its own
// path for them is the generic SpMV (src/util/sparse_matrix.h:72-124).
//
// Structure (persistent CTAs over 128-row tiles, 384 threads, no cluster, no multicast):
//   warp 0   TMA producer: cp.async.bulk.tensor loads of the A tile (128 x 64 bf16) and the B
//            tile (N x 64 bf16) into a 4-stage shared-memory ring, 128-byte swizzle; both tensor
//            maps describe buffers of THIS GPU only
//   warp 1   MMA issuer: one elected lane issues tcgen05.mma (cta_group::1, kind::f16, M=128,
//            N=Npad, K=16) four times per stage, commits the stage back to the producer and, after
//            the last k-block, the accumulator to the epilogue
//   warp 2   allocates / frees the TMEM columns of the two accumulators
//   warps 4-11 epilogue (two per TMEM lane quarter): tcgen05.ld (32 lanes x 32 bit, 16 columns at a time) -> registers ->
//            margins to global memory (128-bit stores) + per-model logistic loss (transposing
//            warp reduction into shared memory, one atomic per CTA and model at the very end)
// The kernel is persistent and the accumulator double-buffered, so the epilogue of one row tile
// overlaps the loads and MMAs of the next.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "../../include/psb200.h"

namespace psb {

constexpr int kTileM = 128, kTileK = 64, kStages = 4, kUmmaK = 16;
constexpr int kDenseThreads = 384;  // 4 service warps + 8 epilogue warps

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {  // arrives on `bar` when all prior MMAs are done
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile in shared memory, rows of 128 bytes, 128-byte swizzle: 8-row groups are
// 1024 bytes apart (stride byte offset), the leading offset is unused for swizzled K-major tiles
__device__ __forceinline__ uint64_t smem_desc_sw128(const void* tile) {
  const uint64_t addr = (uint64_t)(smem_u32(tile) & 0x3ffff) >> 4;
  return addr | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: D fp32, A and B bf16, both K-major, M = 128, N = n
__device__ __forceinline__ uint32_t instr_desc_bf16(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Persistent: CTA c works on row tiles c, c + grid, ...  The smem ring runs across tile
// boundaries, and the accumulator is double-buffered in TMEM (2 x NPAD columns) so that the
// epilogue of tile i overlaps the loads and MMAs of tile i + 1.
template <int NPAD>
__global__ void __launch_bounds__(kDenseThreads, 1)
k_dense_margins(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                float* __restrict__ Y, const float* __restrict__ label, double* __restrict__ loss,
                int n, int d, int n_models) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // 1024-byte aligned tiles (the swizzle atom), then the barriers
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int kBytesA = kTileM * kTileK * 2, kBytesB = NPAD * kTileK * 2;
  unsigned char* tile_a = base;
  unsigned char* tile_b = base + kStages * kBytesA;
  uint64_t* full = reinterpret_cast<uint64_t*>(tile_b + kStages * kBytesB);
  uint64_t* empty = full + kStages;
  uint64_t* acc_full = empty + kStages;   // [2]
  uint64_t* acc_empty = acc_full + 2;     // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  float* loss_s = reinterpret_cast<float*>(tmem_slot + 4);  // [4 epilogue warps][NPAD]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_blocks = (d + kTileK - 1) / kTileK;
  const int n_tiles = (n + kTileM - 1) / kTileM;
  constexpr uint32_t kTmemCols = 2 * NPAD < 32 ? 32 : 2 * NPAD;  // two accumulators, a power of two >= 32

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(acc_full + s, 1); mbar_init(acc_empty + s, 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  } else if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (int i = threadIdx.x; i < 4 * NPAD; i += kDenseThreads) loss_s[i] = 0.f;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ---- TMA producer ----
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
      for (int kb = 0; kb < k_blocks; ++kb, ++it) {
        const int s = it % kStages;
        if (it >= kStages) mbar_wait(empty + s, ((it / kStages) - 1) & 1);
        mbar_expect_tx(full + s, kBytesA + kBytesB);
        tma_load_2d(tile_a + s * kBytesA, &map_x, full + s, kb * kTileK, tile * kTileM);  // rows past n / columns past d read as 0
        tma_load_2d(tile_b + s * kBytesB, &map_w, full + s, kb * kTileK, 0);
      }
  } else if (warp == 1 && lane == 0) {
    // ---- MMA issuer ----
    const uint32_t idesc = instr_desc_bf16(NPAD);
    int it = 0, i = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++i) {
      const int as = i & 1;
      if (i >= 2) {  // the epilogue has drained this accumulator
        mbar_wait(acc_empty + as, ((i >> 1) - 1) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      }
      const uint32_t tmem_acc = tmem_base + (uint32_t)(as * NPAD);
      for (int kb = 0; kb < k_blocks; ++kb, ++it) {
        const int s = it % kStages;
        mbar_wait(full + s, (it / kStages) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint64_t da = smem_desc_sw128(tile_a + s * kBytesA), db = smem_desc_sw128(tile_b + s * kBytesB);
#pragma unroll
        for (int k = 0; k < kTileK / kUmmaK; ++k)  // 16 bf16 = 32 bytes = 2 descriptor units along the swizzled row
          umma_bf16(tmem_acc, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
        umma_commit(empty + s);               // the stage may be refilled once these MMAs have read it
      }
      umma_commit(acc_full + as);
    }
  } else if (warp >= 4) {
    // ---- epilogue: this warp owns TMEM lanes [32 * (warp % 4), +32) = rows of the tile ----
    // two warps per lane quarter: they split the 16-column chunks between them
    const int quarter = warp & 3, half = (warp - 4) >> 2;
    float* my_loss = loss_s + quarter * NPAD;
    const bool vec = (n_models & 3) == 0;
    int i = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++i) {
      const int as = i & 1;
      mbar_wait(acc_full + as, (i >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int row = tile * kTileM + quarter * 32 + lane;
      const float yl = row < n ? label[row] : 0.f;
#pragma unroll
      for (int c0 = half * 16; c0 < NPAD; c0 += 32) {
        if (c0 >= n_models) break;  // uniform
        uint32_t v[16];
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(as * NPAD + c0);
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
            : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        // margins: 16 consecutive floats of this thread's row
        if (Y && row < n) {
          float* dst = Y + (int64_t)row * n_models + c0;
          if (vec && c0 + 16 <= n_models) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
              *reinterpret_cast<float4*>(dst + 4 * q) = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]),
                                                                     __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]));
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c)
              if (c0 + c < n_models) dst[c] = __uint_as_float(v[c]);
          }
        }
        if (loss) {
          // log(1 + exp(-y m)) without overflow; rows past n contribute 0
          float l[16];
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            const float ym = yl * __uint_as_float(v[c]);
            l[c] = row < n ? fmaxf(-ym, 0.f) + __logf(1.f + __expf(-fabsf(ym))) : 0.f;
          }
          // sum over the 32 rows of the warp for 16 models at once: halve the model set while
          // folding lanes (16 + 8 + 4 + 2 + 1 shuffles instead of 16 x 5)
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const float send = (lane & 16) ? l[c] : l[c + 8];
            const float keep = (lane & 16) ? l[c + 8] : l[c];
            l[c] = keep + __shfl_xor_sync(FULL, send, 16);
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float send = (lane & 8) ? l[c] : l[c + 4];
            const float keep = (lane & 8) ? l[c + 4] : l[c];
            l[c] = keep + __shfl_xor_sync(FULL, send, 8);
          }
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const float send = (lane & 4) ? l[c] : l[c + 2];
            const float keep = (lane & 4) ? l[c + 2] : l[c];
            l[c] = keep + __shfl_xor_sync(FULL, send, 4);
          }
          {
            const float send = (lane & 2) ? l[0] : l[1];
            const float keep = (lane & 2) ? l[1] : l[0];
            l[0] = keep + __shfl_xor_sync(FULL, send, 2);
          }
          l[0] += __shfl_xor_sync(FULL, l[0], 1);
          // lane L now holds the sum of model c0 + 8*b4 + 4*b3 + 2*b2 + b1 (bits of L), twice (b0)
          if (!(lane & 1)) {
            const int model = c0 + ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
            my_loss[model] += l[0];  // one lane per model and warp: no atomics
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty + as);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols));
  if (loss)
    for (int m = threadIdx.x; m < n_models; m += kDenseThreads)
      atomicAdd(loss + m, (double)loss_s[m] + (double)loss_s[NPAD + m] + (double)loss_s[2 * NPAD + m] + (double)loss_s[3 * NPAD + m]);
}

// cuTensorMapEncodeTiled through the runtime's driver entry point: no link-time dependency on libcuda
typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiled encode_fn() {
  static EncodeTiled fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess) p = nullptr;
    return (EncodeTiled)p;
  }();
  return fn;
}
// a [rows x cols] bf16 row-major matrix of THIS device, boxes of [box_rows x 64 columns], 128-byte swizzle
static int make_map(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  EncodeTiled enc = encode_fn();
  PSB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kTileK, (cuuint32_t)box_rows};
  const cuuint32_t elem[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, elem,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PSB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d): pointer 16-byte aligned? leading dimension a multiple of 8?", (int)r);
  return 0;
}

template <int NPAD>
static int launch_dense(const CUtensorMap& mx, const CUtensorMap& mw, float* Y, const float* label, double* loss,
                        int n, int d, int n_models, cudaStream_t stream) {
  const size_t smem = 1024 + kStages * (kTileM * kTileK * 2 + NPAD * kTileK * 2) + (2 * kStages + 4) * 8 + 16 + 4 * NPAD * 4;
  PSB_CUDA(cudaFuncSetAttribute(k_dense_margins<NPAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int tiles = (n + kTileM - 1) / kTileM;
  // persistent: as many CTAs as fit (shared memory bound), each walking tiles with stride grid
  int per_sm = 1;
  PSB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_dense_margins<NPAD>, kDenseThreads, smem));
  // every resident CTA holds 2 * NPAD TMEM columns of the 512 of its SM
  const int tmem_fit = 512 / (2 * NPAD < 32 ? 32 : 2 * NPAD);
  if (per_sm > tmem_fit) per_sm = tmem_fit;
  if (per_sm < 1) per_sm = 1;
  const int grid = tiles < per_sm * sm_count() ? tiles : per_sm * sm_count();
  k_dense_margins<NPAD><<<grid, kDenseThreads, smem, stream>>>(mx, mw, Y, label, loss, n, d, n_models);
  PSB_LAUNCH_CHECK();
  return 0;
}

}  // namespace psb

using namespace psb;

// Y[n x n_models] (fp32) = X[n x d] * W[n_models x d]^T with bf16 operands (ld = elements per row,
// multiples of 8; 16-byte aligned pointers); loss[m] += sum_i log(1 + exp(-label_i * Y[i][m])) when
// loss != NULL (zero it first); Y may be NULL when only the losses are wanted.  n_models <= 256.  X, W, Y, label, loss are buffers of the current
// device; the tensor maps built here never describe another GPU's memory.
PSB_API int psb_dense_margins_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, float* Y,
                                   const float* label, double* loss, int32_t n, int32_t d, int32_t n_models,
                                   void* stream) {
  PSB_REQUIRE(X && W && label && (Y || loss) && n >= 1 && d >= 1, "psb_dense_margins_bf16: bad argument");
  PSB_REQUIRE(n_models >= 1 && n_models <= 256, "psb_dense_margins_bf16: n_models=%d outside [1,256]", n_models);
  PSB_REQUIRE(ldx % 8 == 0 && ldw % 8 == 0 && ldx >= d && ldw >= d, "psb_dense_margins_bf16: leading dimensions must be multiples of 8 and >= d");
  int npad = 16;
  while (npad < n_models) npad <<= 1;
  CUtensorMap mx, mw;
  if (int rc = make_map(&mx, X, n, d, ldx, kTileM)) return rc;
  if (int rc = make_map(&mw, W, n_models, d, ldw, npad)) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  switch (npad) {
    case 16: return launch_dense<16>(mx, mw, Y, label, loss, n, d, n_models, s);
    case 32: return launch_dense<32>(mx, mw, Y, label, loss, n, d, n_models, s);
    case 64: return launch_dense<64>(mx, mw, Y, label, loss, n, d, n_models, s);
    case 128: return launch_dense<128>(mx, mw, Y, label, loss, n, d, n_models, s);
    default: return launch_dense<256>(mx, mw, Y, label, loss, n, d, n_models, s);
  }
}
