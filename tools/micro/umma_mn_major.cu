// Microtest: tcgen05.mma kind::f16 with an MN-MAJOR B operand (128-byte swizzle) -- the layout a TMA box of the
// K-major base-normal store [column i][sample] lands in when the SAMPLING GEMM (contraction over columns,
// N = samples) reads it: rows of 128 B = 64 consecutive samples (N), one row per contraction index k.
// One CTA, C (128 x 256) = A (128 x 64, K-major) * B^T with B (256 x 64) held MN-major:
//   byte(n, k) = (n / 64) * LBO + (k / 8) * SBO + (k % 8) * 128 + (((n % 64) / 8) ^ (k % 8)) * 16 + (n % 8) * 2
// The descriptor encodings (LBO, SBO) are passed in; the program reports which candidates reproduce the
// CPU product exactly.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o umma_mn_major umma_mn_major.cu && ./umma_mn_major
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t desc_k_major(uint32_t addr) {      // as tc_gemm.cu make_desc
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ uint64_t desc_mn_major(uint32_t addr, uint32_t lbo16, uint32_t sbo16) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)(lbo16 & 0x3FFF) << 16;
  d |= (uint64_t)(sbo16 & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__global__ void __launch_bounds__(128, 1)
mn_kernel(const __half* __restrict__ A, const __half* __restrict__ B, float* __restrict__ C, uint32_t lbo16,
          uint32_t sbo16, uint32_t bmajor_bit) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* sm = raw + (base - smem_u32(raw));
  uint8_t* sa = sm;                 // 128 rows x 128 B
  uint8_t* sb = sm + 16384;         // 4 chunks x 64 k-rows x 128 B = 32 KB
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_addr;
  for (int e = threadIdx.x; e < 128 * 64; e += 128) {
    const int r = e / 64, k = e % 64;
    *reinterpret_cast<__half*>(sa + r * 128 + (((k / 8) ^ (r % 8)) * 16) + (k % 8) * 2) = A[r * 64 + k];
  }
  for (int e = threadIdx.x; e < 256 * 64; e += 128) {
    const int n = e / 64, k = e % 64;
    const int nn = n % 64;
    *reinterpret_cast<__half*>(sb + (n / 64) * 8192 + k * 128 + (((nn / 8) ^ (k % 8)) * 16) + (nn % 8) * 2) =
        B[n * 64 + k];
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_addr)),
                 "r"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_addr;
  if (threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | (bmajor_bit << 16) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    for (int ks = 0; ks < 4; ++ks) {
      const uint64_t da = desc_k_major(base + ks * 32);
      const uint64_t db = desc_mn_major(base + 16384 + ks * 2048, lbo16, sbo16);
      const uint32_t acc = ks > 0;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc),
          "r"(acc) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar))
                 : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c0 = 0; c0 < 256; c0 += 8) {
    uint32_t v[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(tmem + ((uint32_t)(warp * 32) << 16) + c0) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int u = 0; u < 8; ++u) C[(warp * 32 + lane) * 256 + c0 + u] = __uint_as_float(v[u]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256) : "memory");
}

int main() {
  std::vector<__half> A(128 * 64), B(256 * 64);
  std::vector<float> Af(128 * 64), Bf(256 * 64), ref(128 * 256, 0.f), C(128 * 256);
  uint32_t s = 12345;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (int)((s >> 24) % 17) - 8; };
  for (size_t i = 0; i < A.size(); ++i) { Af[i] = (float)rnd(); A[i] = __float2half(Af[i]); }
  for (size_t i = 0; i < B.size(); ++i) { Bf[i] = (float)rnd(); B[i] = __float2half(Bf[i]); }
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < 256; ++n) {
      float acc = 0.f;
      for (int k = 0; k < 64; ++k) acc += Af[m * 64 + k] * Bf[n * 64 + k];
      ref[m * 256 + n] = acc;
    }
  __half *dA, *dB;
  float* dC;
  cudaMalloc(&dA, A.size() * 2);
  cudaMalloc(&dB, B.size() * 2);
  cudaMalloc(&dC, C.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
  const int smem = 16384 + 32768 + 1024;
  cudaFuncSetAttribute(mn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const uint32_t cand[4][3] = {{8192 >> 4, 1024 >> 4, 1}, {1024 >> 4, 8192 >> 4, 1}, {8192 >> 4, 1024 >> 4, 0},
                               {1, 1024 >> 4, 1}};
  int ok_any = 0;
  for (int c = 0; c < 4; ++c) {
    cudaMemset(dC, 0, C.size() * 4);
    mn_kernel<<<1, 128, smem>>>(dA, dB, dC, cand[c][0], cand[c][1], cand[c][2]);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("candidate %d: CUDA error %s\n", c, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(C.data(), dC, C.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (size_t i = 0; i < C.size(); ++i) bad += C[i] != ref[i];
    printf("candidate %d (LBO=%u SBO=%u b_major=%u): %d mismatches of %zu\n", c, cand[c][0] << 4, cand[c][1] << 4,
           cand[c][2], bad, C.size());
    ok_any |= (bad == 0) << c;
  }
  printf("UMMA_MN_RESULT mask=%d\n", ok_any);
  return 0;
}
