// What one Darlin step costs in random fp64 traffic alone: W warps, each with 4 independent
// 8-byte gathers per lane in flight (then a dependent use), and the same shape as fp64 RED.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/micro/gather_red benchmarks/micro/gather_red.cu
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }
__global__ void k_gather(const double* __restrict__ z, uint32_t n, int rounds, int active_warps, double* out) {
  const int gw = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
  if (gw >= active_warps) return;
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  double acc = 0;
  for (int r = 0; r < rounds; ++r) {
    double v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { s = mix(s + k + (uint32_t)(acc != 12345.678)); v[k] = __ldcg(z + s % n); }
    acc += v[0] + v[1] + v[2] + v[3];  // dependent use: the next round waits for this one
  }
  if (acc == 42.0) out[0] = acc;
}
__global__ void k_red(double* z, uint32_t n, int rounds, int active_warps) {
  const int gw = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
  if (gw >= active_warps) return;
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 999u;
  for (int r = 0; r < rounds; ++r) {
#pragma unroll
    for (int k = 0; k < 4; ++k) { s = mix(s + k); atomicAdd(z + s % n, 1e-9); }
    __threadfence();  // as the barrier arrive does: wait until the reds are performed
  }
}
__global__ void k_mix(double* z, uint32_t n, int rounds, int active_warps, double* out) {
  const int gw = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
  if (gw >= active_warps) return;
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 777u;
  double acc = 0;
  for (int r = 0; r < rounds; ++r) {
    double v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { s = mix(s + k + (uint32_t)(acc != 12345.678)); v[k] = __ldcg(z + s % n); }
    acc += v[0] + v[1] + v[2] + v[3];
#pragma unroll
    for (int k = 0; k < 4; ++k) { s = mix(s + k); atomicAdd(z + s % n, acc * 1e-30); }
    __threadfence();
  }
  if (acc == 42.0) out[0] = acc;
}
int main() {
  const uint32_t n = 677399;
  double *z, *out; cudaMalloc(&z, n * 8); cudaMalloc(&out, 8); cudaMemset(z, 0, n * 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int rounds = 200;
  for (int warps : {148, 296, 592, 1300, 2368}) {
    float tg, tr;
    k_gather<<<296, 256>>>(z, n, 10, warps, out);
    cudaEventRecord(e0); k_gather<<<296, 256>>>(z, n, rounds, warps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&tg, e0, e1);
    k_red<<<296, 256>>>(z, n, 10, warps);
    cudaEventRecord(e0); k_red<<<296, 256>>>(z, n, rounds, warps); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&tr, e0, e1);
    float tm;
    k_mix<<<296, 256>>>(z, n, 10, warps, out);
    cudaEventRecord(e0); k_mix<<<296, 256>>>(z, n, rounds, warps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&tm, e0, e1);
    printf("  gather+red on the same array, one round %.2f us\n", tm * 1e3 / rounds);
    printf("active warps %4d: gather round (128 x 8B per warp) %.2f us = %.1f sectors/ns | red round + fence %.2f us = %.1f red/ns\n",
           warps, tg * 1e3 / rounds, warps * 128.0 / (tg * 1e6 / rounds), tr * 1e3 / rounds, warps * 128.0 / (tr * 1e6 / rounds));
  }
  printf("status %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
