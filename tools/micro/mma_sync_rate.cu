// Microbenchmark: legacy warp-level mma.sync.m16n8k16 (fp16 in, fp32 accumulate) issue rate on sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_sync_rate mma_sync_rate.cu && ./mma_sync_rate
#include <cstdio>
#include <cuda_runtime.h>
#include <cstdint>

__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int NACC>
__global__ void __launch_bounds__(512) rate_kernel(float* out, int iters, uint32_t seed) {
  float acc[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
  uint32_t a[4] = {seed + threadIdx.x, seed * 3u, seed * 5u + threadIdx.x, seed * 7u};
  uint32_t b[2] = {seed * 11u, seed * 13u + threadIdx.x};
  for (int a_ = 0; a_ < 4; ++a_) a[a_] = (a[a_] & 0x03ff03ffu) | 0x3c003c00u;  // halves in [1,2)
  for (int b_ = 0; b_ < 2; ++b_) b[b_] = (b[b_] & 0x03ff03ffu) | 0x3c003c00u;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) mma16816(acc[i], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  float* out;
  cudaMalloc(&out, 148 * 4 * 512 * sizeof(float));
  const int iters = 20000;
  for (int warps = 4; warps <= 16; warps *= 2) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0);
      rate_kernel<8><<<148 * 2, warps * 32>>>(out, iters, 12345u);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double flops = 148.0 * 2 * warps * (double)iters * 8 * 2.0 * 16 * 8 * 16;
      if (rep) printf("warps/CTA %2d (x2 CTA/SM): %.3f ms  %.1f TFLOP/s\n", warps, ms, flops / ms / 1e9);
    }
  }
  return 0;
}
