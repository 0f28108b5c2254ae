// Measures fp64 FMA latency / throughput and the cost of exp() + reciprocal on one SM-ful of warps.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/micro/fp64_rate benchmarks/micro/fp64_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void k_dfma(double* out, int iters, double a, double b) {
  double x[ILP];
  for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x * 1e-3 + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], a, b);
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[gridDim.x * blockDim.x] = (double)(t1 - t0);
}
template <int ILP>
__global__ void k_exp(double* out, int iters, double a) {
  double x[ILP];
  for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x * 1e-3 + i * 0.1;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = 1.0 / (1.0 + exp(x[i] * a));
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[gridDim.x * blockDim.x] = (double)(t1 - t0);
}
int main() {
  double* d; cudaMalloc(&d, (148 * 1024 + 1) * sizeof(double));
  const int iters = 4096;
  for (int warps : {1, 4, 8, 16, 32}) {
    double c1, c4, e1, e4;
    k_dfma<1><<<1, warps * 32>>>(d, iters, 1.0000001, 1e-9); cudaMemcpy(&c1, d + warps * 32, 8, cudaMemcpyDeviceToHost);
    k_dfma<4><<<1, warps * 32>>>(d, iters, 1.0000001, 1e-9); cudaMemcpy(&c4, d + warps * 32, 8, cudaMemcpyDeviceToHost);
    k_exp<1><<<1, warps * 32>>>(d, iters / 16, 0.5); cudaMemcpy(&e1, d + warps * 32, 8, cudaMemcpyDeviceToHost);
    k_exp<4><<<1, warps * 32>>>(d, iters / 16, 0.5); cudaMemcpy(&e4, d + warps * 32, 8, cudaMemcpyDeviceToHost);
    printf("warps/SM %2d: dfma ilp1 %.1f cyc/instr (per warp), ilp4 %.1f; lanes/clk/SM ilp4 %.1f | sigmoid(exp+div) ilp1 %.0f cyc, ilp4 %.0f cyc each; lanes/clk/SM %.2f\n",
           warps, c1 / iters, c4 / iters / 4, warps * 32.0 * 4 * iters / c4, e1 / (iters / 16), e4 / (iters / 16) / 4,
           warps * 32.0 * 4 * (iters / 16) / e4);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status %s\n", cudaGetErrorString(e));
  return 0;
}
