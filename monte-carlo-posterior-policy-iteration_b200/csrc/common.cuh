// Shared device/host helpers for the PPI update-path kernels (sm_100a only).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#include "../../include/ppi_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "ppi_b200 kernels are written for sm_100a (B200) only"
#endif

namespace ppi {

void set_error(const char* fmt, ...);

void count_launch(int n);

// Records a CUDA-event pair around the launches in its scope when ppi_profile_enable(1) is set.
class ProfileScope {
 public:
  ProfileScope(int slot, cudaStream_t stream);
  ~ProfileScope();
 private:
  int slot_;
  cudaStream_t stream_;
  bool on_;
  size_t index_ = 0;
};

// `n` = kernels launched since the previous check
inline int check_launch(const char* what, int n = 1) {
  count_launch(n);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

#define PPI_CUDA_TRY(expr)                                                  \
  do {                                                                      \
    cudaError_t _e = (expr);                                                \
    if (_e != cudaSuccess) {                                                \
      ppi::set_error("%s: %s", #expr, cudaGetErrorString(_e));              \
      return (int)_e;                                                       \
    }                                                                       \
  } while (0)

#define PPI_REQUIRE(cond, code, ...)   \
  do {                                 \
    if (!(cond)) {                     \
      ppi::set_error(__VA_ARGS__);     \
      return (code);                   \
    }                                  \
  } while (0)

constexpr int kNumSMs = 148;  // B200

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011).  counter = (elem_quad lo, elem_quad hi, draw lo, draw hi),
// key = seed.  One call -> four uint32 -> four standard normals (two Box-Muller pairs).
// ---------------------------------------------------------------------------------------------
struct Philox4 {
  uint32_t x, y, z, w;
};

__device__ __forceinline__ Philox4 philox4x32_10(uint64_t ctr_lo, uint64_t ctr_hi, uint64_t seed) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  constexpr uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32);
  uint32_t c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  return Philox4{c0, c1, c2, c3};
}

// Same function with the 10 round keys precomputed (they depend on the seed only): hot loops that
// draw many quads per thread hoist the key schedule out of the loop (20 registers, 20 fewer
// instructions per call).
struct PhiloxKeys {
  uint32_t k0[10], k1[10];
};
__device__ __forceinline__ PhiloxKeys philox_keys(uint64_t seed) {
  constexpr uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  PhiloxKeys ks;
  uint32_t a = (uint32_t)seed, b = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    ks.k0[r] = a;
    ks.k1[r] = b;
    a += W0;
    b += W1;
  }
  return ks;
}
__device__ __forceinline__ Philox4 philox4x32_10(uint64_t ctr_lo, uint64_t ctr_hi, const PhiloxKeys& ks) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32);
  uint32_t c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    const uint32_t n0 = hi1 ^ c1 ^ ks.k0[r], n2 = hi0 ^ c3 ^ ks.k1[r];
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
  }
  return Philox4{c0, c1, c2, c3};
}

// (0,1] uniform from 32 random bits; never 0 so that log() is finite.
__device__ __forceinline__ float u01(uint32_t r) {
  return fmaf((float)r, 2.3283064365386963e-10f, 1.1641532182693481e-10f);
}

// Box-Muller.  The base normals are emitted as IEEE half-precision VALUES (held in fp32): 11
// significant bits, relative spacing 2^-11 (variance inflation < 2^-22/12, far below the error of
// __logf / __sincosf), |z| <= 6.67 so nothing overflows.  A Philox normal is therefore an EXACT
// tensor-core operand (fp16, and also tf32): the dense Gaussian transform z L^T needs no low-order
// part of z (tc_gemm.cu).  Every consumer (ppi_philox_normal and the fused samplers) goes through
// this function, so they all see the same numbers.
__device__ __forceinline__ float round_half(float v) { return __half2float(__float2half_rn(v)); }

// [0, 1) from the top 23 bits without an integer -> float conversion (those share the 16-lane XU pipe
// with the MUFU transcendentals, which is what bounds the generator): the angle needs no more.
__device__ __forceinline__ float u01_23(uint32_t r) { return __uint_as_float(0x3F800000u | (r >> 9)) - 1.0f; }

// Radius and direction of one Box-Muller pair: r = sqrt(-2 ln u), (c, s) = (cos, sin)(2 pi v), v the top 23 bits
// of b.  Written as the bare MUFU forms: lg2.approx.ftz (u >= 2^-33 is never subnormal -- the non-ftz form
// costs a compare and two predicated fix-ups per call), and the angle as 2 pi t with t in [1, 2) read straight
// off the bit pattern (sine and cosine have period one in t; the range reduction inside MUFU keeps the
// fractional part, 23 bits either way).
__device__ __forceinline__ void box_muller_polar(uint32_t a, uint32_t b, float& r, float& s, float& c) {
  const float u = u01(a);
  const float t = __uint_as_float(0x3F800000u | (b >> 9));
  float lg;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(u));
  // the radicand is clamped from below: MUFU.LG2 carries an absolute error of 2^-22, so for u in the top
  // float bins below 1 nothing guarantees a non-positive logarithm (sqrt of a negative number would be a NaN
  // that poisons a whole D x D moment), and u01() returns exactly 1 for r >= 2^32 - 128 (an atom at z = 0)
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaxf(-1.3862943611198906f * lg, 1e-12f)));  // sqrt(-2 ln u)
  const float x = 6.283185307179586f * t;
  asm("sin.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(x));
  asm("cos.approx.ftz.f32 %0, %1;" : "=f"(c) : "f"(x));
}

// One Box-Muller pair as a packed half2 (z0 low, z1 high) -- the form the tensor-core operand
// stores want; box_muller() below widens exactly these values.
__device__ __forceinline__ __half2 box_muller_h2(uint32_t a, uint32_t b) {
  // every transcendental is a single MUFU instruction (lg2 / sqrt / sin / cos, no slow paths): the
  // result is rounded to 11 bits anyway
  float r, s, c;
  box_muller_polar(a, b, r, s, c);
  return __floats2half2_rn(r * c, r * s);
}

// The same pair WITHOUT the rounding to the fp16 grid (ppi_philox_normal_full): what the generator
// would emit if the normals were not made exact tensor-core operands.  Same Philox words, same angles.
__device__ __forceinline__ void box_muller_full(uint32_t a, uint32_t b, float& z0, float& z1) {
  float r, s, c;
  box_muller_polar(a, b, r, s, c);
  z0 = r * c;
  z1 = r * s;
}

__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& z0, float& z1) {
  const float2 z = __half22float2(box_muller_h2(a, b));
  z0 = z.x;
  z1 = z.y;
}

// The four normals of element quad q (elements 4q..4q+3 of the row-major global (K_total, n)
// base-normal matrix) for draw `draw_id`.
__device__ __forceinline__ void philox_normal4(uint64_t seed, uint64_t draw_id, uint64_t quad,
                                               float out[4]) {
  Philox4 p = philox4x32_10(quad, draw_id, seed);
  box_muller(p.x, p.y, out[0], out[1]);
  box_muller(p.z, p.w, out[2], out[3]);
}

__device__ __forceinline__ void philox_normal4(const PhiloxKeys& ks, uint64_t draw_id, uint64_t quad,
                                               float out[4]) {
  Philox4 p = philox4x32_10(quad, draw_id, ks);
  box_muller(p.x, p.y, out[0], out[1]);
  box_muller(p.z, p.w, out[2], out[3]);
}

// Same four normals as two packed half2 (elements 4q, 4q+1 | 4q+2, 4q+3).
__device__ __forceinline__ void philox_normal4_h2(const PhiloxKeys& ks, uint64_t draw_id, uint64_t quad,
                                                  __half2& lo, __half2& hi) {
  Philox4 p = philox4x32_10(quad, draw_id, ks);
  lo = box_muller_h2(p.x, p.y);
  hi = box_muller_h2(p.z, p.w);
}

// ---- column-quad order (the dense Gaussian sampler, D > 128) ------------------------------------------------
// The tensor-core sampler keeps its base normals K-major (tc_gemm.cu: zT[column][sample]), so its Philox
// counters run DOWN the columns: one call yields the normals of samples 4Q .. 4Q+3 of ONE column, i.e. 8 bytes
// that are contiguous in that store (no transposition).  counter = (Q, draw << 32 | column), key = seed ^ tag
// (the tag keeps the stream apart from the row-quad stream of the same seed).  Element (row, col) is still a
// pure function of (seed, draw_id, global row, col): K-sharded ranks reproduce the 1-GPU matrix.
constexpr uint64_t kPhiloxColTag = 0x9E3779B97F4A7C15ull;

__device__ __forceinline__ void philox_cq_normal4_h2(const PhiloxKeys& ks_tagged, uint64_t draw_id, uint32_t col,
                                                     uint64_t rowquad, __half2& lo, __half2& hi) {
  const Philox4 p = philox4x32_10(rowquad, (draw_id << 32) | (uint64_t)col, ks_tagged);
  lo = box_muller_h2(p.x, p.y);
  hi = box_muller_h2(p.z, p.w);
}

// one element (slow path: unaligned shards, the TF32 scheme, host-visible copies)
__device__ __forceinline__ float philox_cq_normal1(uint64_t seed, uint64_t draw_id, uint64_t row, uint32_t col) {
  const Philox4 p = philox4x32_10(row >> 2, (draw_id << 32) | (uint64_t)col, seed ^ kPhiloxColTag);
  float z0, z1;
  if ((row & 2) == 0) box_muller(p.x, p.y, z0, z1);
  else box_muller(p.z, p.w, z0, z1);
  return (row & 1) ? z1 : z0;
}

// Single element e of the global matrix (slow path for unaligned access).
__device__ __forceinline__ float philox_normal1(uint64_t seed, uint64_t draw_id, uint64_t elem) {
  float v[4];
  philox_normal4(seed, draw_id, elem >> 2, v);
  return v[elem & 3];
}

// ---------------------------------------------------------------------------------------------
// warp / block reductions
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Sum over the block; result valid in every thread.  `scratch` holds >= 32 T.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  T r = (lane < nw) ? scratch[lane] : T(0);
  r = warp_sum(r);
  return r;
}

__device__ __forceinline__ float clipf(float v, float lo, float hi) {
  // np.clip(x, a_min, a_max) = minimum(maximum(x, a_min), a_max); NaN propagates.
  float r;
  asm("max.NaN.f32 %0, %1, %2;\n\tmin.NaN.f32 %0, %0, %3;" : "=&f"(r) : "f"(v), "f"(lo), "f"(hi));
  return r;
}

// In-place right-looking Cholesky of the n x n matrix in shared memory `a` (lower part used).
// Returns through *fail the 1-based index of the first non-positive pivot (0 = success).
__device__ inline void block_cholesky(double* a, int n, int* fail_smem) {
  // Right-looking, ONE barrier per column: the trailing update uses the unscaled column j,
  // a[i][k] -= a[i][j] a[k][j] / a[j][j]; column j is final after step j and is scaled by
  // 1 / sqrt(pivot) in a last pass.  Every thread reads the same pivot, so the failure exit is uniform.
  int fail = 0;
  __syncthreads();
  for (int j = 0; j < n; ++j) {
    const double p = a[j * n + j];
    if (!(p > 0.0) || !isfinite(p)) {
      fail = j + 1;
      break;
    }
    const double inv = 1.0 / p;
    const int m = n - j - 1;  // trailing size
    for (int e = threadIdx.x; e < m * m; e += blockDim.x) {
      const int r = e / m, c = e - r * m;
      if (c <= r) {
        const int i = j + 1 + r, k = j + 1 + c;
        a[i * n + k] -= a[i * n + j] * a[k * n + j] * inv;
      }
    }
    __syncthreads();
  }
  const int done = fail ? fail - 1 : n;  // columns with a valid pivot
  for (int e = threadIdx.x; e < n * done; e += blockDim.x) {
    const int i = e / done, j = e - i * done;
    if (j < i) a[i * n + j] /= sqrt(a[j * n + j]);
  }
  __syncthreads();
  for (int j = threadIdx.x; j < n; j += blockDim.x)
    a[j * n + j] = j < done ? sqrt(a[j * n + j]) : nan("");
  if (threadIdx.x == 0) *fail_smem = fail;
  __syncthreads();
}

}  // namespace ppi
