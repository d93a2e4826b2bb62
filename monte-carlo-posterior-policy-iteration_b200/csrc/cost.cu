// Batched cost evaluation: row-wise reductions (K, D) -> (K,).
//   test functions        optimization/functions.py:21-104
//   tracking cost         policy_search/envs.py:65-72,83-87  (the "synthetic trajectory return")
// One warp per sample row, float4 loads when the row stride allows, fp32 accumulation per lane
// (<= D/32 terms) followed by a shuffle tree.  HBM-bound: 4*D bytes read per sample.
#include "common.cuh"

namespace ppi {

template <int FN>
__device__ __forceinline__ float term(float x) {
  if (FN == PPI_FN_STYBLINSKI) {
    const float x2 = x * x;
    return fmaf(x2, x2, fmaf(-16.f, x2, 5.f * x));
  } else {  // Rastrigin: x^2 - 10 cos(2 pi x).  Exact range reduction (x - rint(x) is exact in fp32)
    // then one MUFU.COS on [-pi, pi] (abs. error ~5e-7): cospif() costs ~40 instructions and made the
    // kernel issue-bound instead of HBM-bound
    const float fr = x - rintf(x);
    return fmaf(x, x, -10.f * __cosf(6.283185307179586f * fr));
  }
}

__device__ __forceinline__ float rosen(float a, float b) {  // a = x[i], b = x[i+1]
  const float u = b - a * a, v = 1.f - a;
  return fmaf(100.f * u, u, v * v);
}

template <int FN>
__global__ void __launch_bounds__(256)
cost_rowwise_kernel(const float* __restrict__ x, int64_t K, int64_t D, float* __restrict__ c) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
  for (int64_t b = warp; b < K; b += nwarps) {
    const float* row = x + b * D;
    float acc = 0.f;
    if (FN == PPI_FN_ROSENBROCK) {
      if (vec) {
        const int64_t nq = D / 4;
        for (int64_t q0 = 0; q0 < nq; q0 += 32) {
          const int64_t q = q0 + lane;
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (q < nq) v = __ldg(reinterpret_cast<const float4*>(row) + q);
          float nxt = __shfl_down_sync(0xffffffffu, v.x, 1);
          if (lane == 31 && q + 1 < nq) nxt = __ldg(row + 4 * (q + 1));
          if (q < nq) {
            acc += rosen(v.x, v.y) + rosen(v.y, v.z) + rosen(v.z, v.w);
            if (q + 1 < nq) acc += rosen(v.w, nxt);
          }
        }
      } else {
        for (int64_t i = lane; i + 1 < D; i += 32) acc += rosen(__ldg(row + i), __ldg(row + i + 1));
      }
    } else {
      if (vec) {
        const int64_t nq = D / 4;
        for (int64_t q = lane; q < nq; q += 32) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(row) + q);
          acc += (term<FN>(v.x) + term<FN>(v.y)) + (term<FN>(v.z) + term<FN>(v.w));
        }
      } else {
        for (int64_t i = lane; i < D; i += 32) acc += term<FN>(__ldg(row + i));
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      if (FN == PPI_FN_STYBLINSKI) acc = fmaf(0.5f, acc, 39.16599f * (float)D);
      if (FN == PPI_FN_RASTRIGIN) acc += 10.f * (float)D;
      c[b] = acc;
    }
  }
}

__global__ void __launch_bounds__(256)
cost_himmelblau_kernel(const float* __restrict__ x, int64_t K, int64_t D, float* __restrict__ c) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < K; b += stride) {
    const float x0 = x[b * D], x1 = x[b * D + 1];
    const float p = x0 * x0 + x1 - 11.f, q = x0 + x1 * x1 - 7.f;
    c[b] = -1.f * (p * p + q * q);
  }
}

// x^T A x + 0.01 noise: one warp per sample, x row staged in shared memory.
__global__ void __launch_bounds__(256)
cost_noisy_sphere_kernel(const float* __restrict__ x, int64_t K, int D, const float* __restrict__ A,
                         const float* __restrict__ noise, float* __restrict__ c) {
  extern __shared__ float sx[];  // 8 warps * D
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* xr = sx + wib * D;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < K; b += nwarps) {
    for (int i = lane; i < D; i += 32) xr[i] = x[b * D + i];
    __syncwarp();
    float acc = 0.f;
    for (int i = lane; i < D; i += 32) {
      const float* ar = A + (int64_t)i * D;
      float y = 0.f;
      for (int j = 0; j < D; ++j) y = fmaf(__ldg(ar + j), xr[j], y);
      acc = fmaf(xr[i], y, acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) c[b] = acc + (noise ? 0.01f * noise[b] : 0.f);
    __syncwarp();
  }
}

// Tracking cost: mean over (t, j < dt) of (x[b,t,j] - target[t,j])^2; one warp per sample.
__global__ void __launch_bounds__(256)
cost_tracking_kernel(const float* __restrict__ x, int64_t K, int T, int dy,
                     const float* __restrict__ target, int dt, float* __restrict__ c) {
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int n = T * dt;
  const float inv = 1.f / (float)n;
  // 16-byte loads need aligned bases and sample rows that are a whole number of 16-byte words
  const bool vec_ok = (((uintptr_t)x | (uintptr_t)target) & 15) == 0 && (((int64_t)T * dy) & 3) == 0;
  for (int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < K; b += nwarps) {
    const float* xb = x + b * (int64_t)T * dy;
    float acc = 0.f;
    if (dy == 4 && dt == 2 && vec_ok) {
      // value + derivative rows (feature policies with use_derivatives): one 16-byte load per time step,
      // the two value channels enter the cost; four time steps in flight per lane
      const float4* x4 = reinterpret_cast<const float4*>(xb);
      const float2* t2 = reinterpret_cast<const float2*>(target);
      float a0 = 0.f, a1 = 0.f;
      int t = lane;
      for (; t + 96 < T; t += 128) {
        float4 v[4];
        float2 g[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          v[u] = __ldg(x4 + t + 32 * u);
          g[u] = __ldg(t2 + t + 32 * u);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float d0 = v[u].x - g[u].x, d1 = v[u].y - g[u].y;
          a0 = fmaf(d0, d0, a0);
          a1 = fmaf(d1, d1, a1);
        }
      }
      for (; t < T; t += 32) {
        const float4 v = __ldg(x4 + t);
        const float2 g = __ldg(t2 + t);
        const float d0 = v.x - g.x, d1 = v.y - g.y;
        a0 = fmaf(d0, d0, a0);
        a1 = fmaf(d1, d1, a1);
      }
      acc = a0 + a1;
    } else if (dt == dy) {
      for (int e = lane; e < n; e += 32) {
        const float df = __ldg(xb + e) - __ldg(target + e);
        acc = fmaf(df, df, acc);
      }
    } else {
      for (int e = lane; e < n; e += 32) {
        const int t = e / dt, j = e % dt;
        const float df = __ldg(xb + t * dy + j) - __ldg(target + e);
        acc = fmaf(df, df, acc);
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) c[b] = acc * inv;
  }
}

}  // namespace ppi

using namespace ppi;

extern "C" int ppi_cost_function(int fn, const float* x, int64_t K, int64_t D, const float* aux,
                                 const float* noise, float* c, void* stream) {
  PPI_REQUIRE(x && c && K >= 0 && D > 0, PPI_ERR_INVALID, "ppi_cost_function: bad arguments");
  if (K == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t warps_needed = K;
  int grid = (int)(ceil_div(warps_needed, 8) < (int64_t)kNumSMs * 8 ? ceil_div(warps_needed, 8)
                                                                   : (int64_t)kNumSMs * 8);
  ProfileScope prof(PPI_PROF_COST, st);
  switch (fn) {
    case PPI_FN_HIMMELBLAU:
      PPI_REQUIRE(D >= 2, PPI_ERR_INVALID, "Himmelblau needs D >= 2");
      grid = (int)(ceil_div(K, 256) < kNumSMs * 8 ? ceil_div(K, 256) : kNumSMs * 8);
      cost_himmelblau_kernel<<<grid, 256, 0, st>>>(x, K, D, c);
      break;
    case PPI_FN_ROSENBROCK:
      cost_rowwise_kernel<PPI_FN_ROSENBROCK><<<grid, 256, 0, st>>>(x, K, D, c);
      break;
    case PPI_FN_STYBLINSKI:
      cost_rowwise_kernel<PPI_FN_STYBLINSKI><<<grid, 256, 0, st>>>(x, K, D, c);
      break;
    case PPI_FN_RASTRIGIN:
      cost_rowwise_kernel<PPI_FN_RASTRIGIN><<<grid, 256, 0, st>>>(x, K, D, c);
      break;
    case PPI_FN_NOISY_SPHERE: {
      PPI_REQUIRE(aux, PPI_ERR_INVALID, "NoisySphere needs the (D,D) matrix A in aux");
      const size_t smem = 8 * (size_t)D * sizeof(float);
      PPI_REQUIRE(smem <= 200 * 1024, PPI_ERR_UNSUPPORTED, "NoisySphere supports D <= 6400");
      if (smem > 48 * 1024)
        PPI_CUDA_TRY(cudaFuncSetAttribute(cost_noisy_sphere_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      cost_noisy_sphere_kernel<<<grid, 256, smem, st>>>(x, K, (int)D, aux, noise, c);
      break;
    }
    default:
      PPI_REQUIRE(false, PPI_ERR_INVALID, "unknown cost function id %d", fn);
  }
  return check_launch("cost kernel");
}

extern "C" int ppi_cost_tracking(const float* x, int64_t K, int T, int dy, const float* target,
                                 int dt, float* c, void* stream) {
  PPI_REQUIRE(x && target && c && K >= 0 && T > 0 && dy > 0 && dt > 0 && dt <= dy, PPI_ERR_INVALID,
              "ppi_cost_tracking: bad arguments");
  if (K == 0) return 0;
  const int grid = (int)(ceil_div(K, 8) < (int64_t)kNumSMs * 8 ? ceil_div(K, 8) : (int64_t)kNumSMs * 8);
  ProfileScope prof(PPI_PROF_COST, (cudaStream_t)stream);
  cost_tracking_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, K, T, dy, target, dt, c);
  return check_launch("cost_tracking_kernel");
}
