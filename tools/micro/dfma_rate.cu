// fp64 FMA issue rate of one SM / the whole B200 (CUDA cores, not DMMA): decides whether the fp64
// single-block helpers (kernel_shift_kernel, block_cholesky) are bound by the fp64 pipe.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dfma_rate dfma_rate.cu && ./dfma_rate
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_kernel(double* out, int iters, double a, double b) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void ffma_kernel(float* out, int iters, float a, float b) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 1e-3f + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  double* d;
  cudaMalloc(&d, 148 * 4 * 1024 * sizeof(double));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 20000;
  for (int blocks : {1, 148, 148 * 2}) {
    for (int threads : {128, 512, 1024}) {
      dfma_kernel<<<blocks, threads>>>(d, 100, 0.999, 1e-3);
      cudaEventRecord(e0);
      dfma_kernel<<<blocks, threads>>>(d, iters, 0.999, 1e-3);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      const double fma = (double)blocks * threads * 16.0 * iters;
      printf("DFMA blocks %4d threads %4d: %8.3f ms  %8.2f GFMA/s total  %7.2f FMA/clk/SM @1.9GHz\n", blocks,
             threads, ms, fma / ms * 1e-6, fma / (ms * 1e-3) / 1.9e9 / (blocks < 148 ? blocks : 148));
      ffma_kernel<<<blocks, threads>>>((float*)d, 100, 0.999f, 1e-3f);
      cudaEventRecord(e0);
      ffma_kernel<<<blocks, threads>>>((float*)d, iters, 0.999f, 1e-3f);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      cudaEventElapsedTime(&ms, e0, e1);
      printf("FFMA blocks %4d threads %4d: %8.3f ms  %8.2f GFMA/s total  %7.2f FMA/clk/SM @1.9GHz\n", blocks,
             threads, ms, fma / ms * 1e-6, fma / (ms * 1e-3) / 1.9e9 / (blocks < 148 ? blocks : 148));
    }
  }
  return 0;
}
