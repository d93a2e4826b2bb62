// Action-prior sampling kernels.
//   philox_normal            np.random.randn                      ppi/samplers.py:18-20
//   gram_chol / chol_small   kernel Gram + np.linalg.cholesky     ppi/policies.py:312,923-1091
//   sample_matnorm           BaseKernel.__call__                  ppi/policies.py:735-745
//   sample_features          BaseFeatures.__call__                ppi/policies.py:404-421
//   sample_iid               WhiteNoiseIid.__call__               ppi/policies.py:1166-1174
//   sample_gaussian (small)  mu + z L^T                           ppi/samplers.py:38-39
// The small matrix products (H, d, F <= ~128) run as fp32 SIMT with shared-memory operands;
// the large dense Gaussian transform lives in tc_gemm.cu (tcgen05).
#include "common.cuh"

namespace ppi {

// ------------------------------------------------------------------------------------------
// Philox normals -> global memory
// ------------------------------------------------------------------------------------------
template <bool FULL>
__global__ void __launch_bounds__(256) philox_normal_kernel(float* __restrict__ out, int64_t n_local,
                                                            uint64_t elem_base, uint64_t seed,
                                                            uint64_t draw) {
  const uint64_t q0 = elem_base >> 2;
  const uint64_t q1 = (elem_base + (uint64_t)n_local + 3) >> 2;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t q = q0 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < q1; q += stride) {
    float v[4];
    if (FULL) {
      const Philox4 p = philox4x32_10(q, draw, seed);
      box_muller_full(p.x, p.y, v[0], v[1]);
      box_muller_full(p.z, p.w, v[2], v[3]);
    } else {
      philox_normal4(seed, draw, q, v);
    }
    const int64_t local = (int64_t)(q * 4) - (int64_t)elem_base;
    if (local >= 0 && local + 4 <= n_local &&
        ((reinterpret_cast<uintptr_t>(out + local) & 15) == 0)) {
      *reinterpret_cast<float4*>(out + local) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t l = local + i;
        if (l >= 0 && l < n_local) out[l] = v[i];
      }
    }
  }
}

// Fill shared memory s[0..n) with base normals elem_base .. elem_base+n (Philox) or copy them
// from the injected matrix z (global element index = elem_base - z_base).
__device__ __forceinline__ void load_normals(float* s, int n, const float* __restrict__ z,
                                             uint64_t z_local_base, uint64_t elem_base,
                                             uint64_t seed, uint64_t draw) {
  if (z != nullptr) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) s[i] = z[z_local_base + i];
  } else {
    const uint64_t q0 = elem_base >> 2;
    const int nq = (int)(((elem_base + n + 3) >> 2) - q0);
    for (int qi = threadIdx.x; qi < nq; qi += blockDim.x) {
      float v[4];
      philox_normal4(seed, draw, q0 + qi, v);
      const int64_t local = (int64_t)((q0 + qi) * 4) - (int64_t)elem_base;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t l = local + i;
        if (l >= 0 && l < n) s[l] = v[i];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// Matrix-normal sampler  x = bias + L_in (z L_out^T), S samples per block iteration.
// Each thread owns one output column l and the row pair (p, H-1-p) for S samples, so every
// L_in element read from shared memory feeds S FMAs and the triangular work is balanced.
// ------------------------------------------------------------------------------------------
template <int S>
__global__ void __launch_bounds__(256)
sample_matnorm_kernel(const float* __restrict__ z, uint64_t seed, uint64_t draw, uint64_t row_offset,
                      int64_t K, int H, int d, const float* __restrict__ mean_fn,
                      const float* __restrict__ mean, const float* __restrict__ l_in,
                      const float* __restrict__ l_out, int l_out_diag,
                      const float* __restrict__ lower, const float* __restrict__ upper,
                      float* __restrict__ x) {
  extern __shared__ float smem[];
  const int HD = H * d;
  float* sLin = smem;               // H*H
  float* sLout = sLin + H * H;      // d*d
  float* sBias = sLout + d * d;     // H*d
  float* sZ = sBias + HD;           // S*H*d
  float* sU = sZ + S * HD;          // S*H*d

  for (int i = threadIdx.x; i < H * H; i += blockDim.x) sLin[i] = l_in[i];
  for (int i = threadIdx.x; i < d * d; i += blockDim.x) sLout[i] = l_out[i];
  for (int i = threadIdx.x; i < HD; i += blockDim.x) sBias[i] = mean_fn[i % d] + mean[i];
  __syncthreads();

  const int64_t n_groups = ceil_div(K, S);
  const int half = (H + 1) / 2;
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t b0 = g * S;
    const int ns = (int)min((int64_t)S, K - b0);
    load_normals(sZ, ns * HD, z, (uint64_t)b0 * HD, (row_offset + (uint64_t)b0) * HD, seed, draw);
    __syncthreads();
    // U = Z L_out^T
    for (int e = threadIdx.x; e < ns * HD; e += blockDim.x) {
      const int l = e % d;
      const float* zr = sZ + (e - l);
      float u;
      if (l_out_diag) {
        u = zr[l] * sLout[l * d + l];
      } else {
        u = 0.f;
        for (int j = 0; j < d; ++j) u = fmaf(zr[j], sLout[l * d + j], u);
      }
      sU[e] = u;
    }
    __syncthreads();
    // X = bias + L_in U
    for (int it = threadIdx.x; it < half * d; it += blockDim.x) {
      const int l = it % d, p = it / d;
      const int k1 = p, k2 = H - 1 - p;  // k1 <= k2
      float a1[S], a2[S];
#pragma unroll
      for (int s = 0; s < S; ++s) a1[s] = a2[s] = 0.f;
      const float* r1 = sLin + k1 * H;
      const float* r2 = sLin + k2 * H;
      for (int i = 0; i <= k1; ++i) {
        const float w1 = r1[i], w2 = r2[i];
#pragma unroll
        for (int s = 0; s < S; ++s) {
          const float u = sU[s * HD + i * d + l];
          a1[s] = fmaf(w1, u, a1[s]);
          a2[s] = fmaf(w2, u, a2[s]);
        }
      }
      for (int i = k1 + 1; i <= k2; ++i) {
        const float w2 = r2[i];
#pragma unroll
        for (int s = 0; s < S; ++s) a2[s] = fmaf(w2, sU[s * HD + i * d + l], a2[s]);
      }
      const float lo = lower ? lower[l] : -INFINITY, hi = upper ? upper[l] : INFINITY;
#pragma unroll
      for (int s = 0; s < S; ++s) {
        if (s < ns) {
          float* xr = x + (b0 + s) * HD;
          xr[k1 * d + l] = clipf(sBias[k1 * d + l] + a1[s], lo, hi);
          if (k2 != k1) xr[k2 * d + l] = clipf(sBias[k2 * d + l] + a2[s], lo, hi);
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// Feature sampler: w = mean + L_in z L_out^T (K,F,d); y = [mean_fn + Phi w, Phi' w] (K,T,dy).
// grid.y tiles the time axis (TT steps); each block recomputes the tiny w for its samples.
// ------------------------------------------------------------------------------------------
constexpr int kFeatTT = 128;

template <int S>
__global__ void __launch_bounds__(256)
sample_features_kernel(const float* __restrict__ z, uint64_t seed, uint64_t draw, uint64_t row_offset,
                       int64_t K, int F, int d, int T, const float* __restrict__ mean_fn,
                       const float* __restrict__ mean, const float* __restrict__ l_in,
                       const float* __restrict__ l_out, const float* __restrict__ feat,
                       const float* __restrict__ dfeat, const float* __restrict__ lower,
                       const float* __restrict__ upper, float* __restrict__ w,
                       float* __restrict__ y) {
  extern __shared__ float smem[];
  const int FD = F * d;
  const int dy = dfeat ? 2 * d : d;
  float* sPhi = smem;                       // F*TT (transposed tile)
  float* sdPhi = sPhi + F * kFeatTT;        // F*TT
  float* sLin = sdPhi + F * kFeatTT;        // F*F
  float* sLout = sLin + F * F;              // d*d
  float* sMean = sLout + d * d;             // F*d
  float* sZ = sMean + FD;                   // S*F*d
  float* sU = sZ + S * FD;                  // S*F*d
  float* sW = sU + S * FD;                  // S*F*d

  const int t0 = blockIdx.y * kFeatTT;
  const int nt = min(kFeatTT, T - t0);
  for (int i = threadIdx.x; i < F * kFeatTT; i += blockDim.x) {
    const int f = i / kFeatTT, tl = i % kFeatTT;
    const bool ok = tl < nt;
    sPhi[i] = ok ? feat[(int64_t)(t0 + tl) * F + f] : 0.f;
    sdPhi[i] = (ok && dfeat) ? dfeat[(int64_t)(t0 + tl) * F + f] : 0.f;
  }
  for (int i = threadIdx.x; i < F * F; i += blockDim.x) sLin[i] = l_in[i];
  for (int i = threadIdx.x; i < d * d; i += blockDim.x) sLout[i] = l_out[i];
  for (int i = threadIdx.x; i < FD; i += blockDim.x) sMean[i] = mean[i];
  __syncthreads();

  const int64_t n_groups = ceil_div(K, S);
  const int tl = threadIdx.x % kFeatTT;
  const int srow = threadIdx.x / kFeatTT, nsrow = blockDim.x / kFeatTT;
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t b0 = g * S;
    const int ns = (int)min((int64_t)S, K - b0);
    load_normals(sZ, ns * FD, z, (uint64_t)b0 * FD, (row_offset + (uint64_t)b0) * FD, seed, draw);
    __syncthreads();
    for (int e = threadIdx.x; e < ns * FD; e += blockDim.x) {
      const int l = e % d;
      const float* zr = sZ + (e - l);
      float u = 0.f;
      for (int j = 0; j < d; ++j) u = fmaf(zr[j], sLout[l * d + j], u);
      sU[e] = u;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < ns * FD; e += blockDim.x) {
      const int s = e / FD, r = e % FD, k = r / d, l = r % d;
      float a = 0.f;
      const float* row = sLin + k * F;
      for (int i = 0; i <= k; ++i) a = fmaf(row[i], sU[s * FD + i * d + l], a);
      a += sMean[r];
      sW[e] = a;
      if (blockIdx.y == 0) w[(b0 + s) * FD + r] = a;
    }
    __syncthreads();
    if (tl < nt) {
      for (int s = srow; s < ns; s += nsrow) {
        float* yr = y + ((b0 + s) * (int64_t)T + (t0 + tl)) * dy;
        for (int c0 = 0; c0 < d; c0 += 4) {
          float ax[4] = {0.f, 0.f, 0.f, 0.f}, ad[4] = {0.f, 0.f, 0.f, 0.f};
          const int nc = min(4, d - c0);
          for (int f = 0; f < F; ++f) {
            const float ph = sPhi[f * kFeatTT + tl], dph = sdPhi[f * kFeatTT + tl];
            const float* wr = sW + s * FD + f * d + c0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (j < nc) {
                ax[j] = fmaf(ph, wr[j], ax[j]);
                ad[j] = fmaf(dph, wr[j], ad[j]);
              }
            }
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (j < nc) {
              const int c = c0 + j;
              const float lo = lower ? lower[c] : -INFINITY, hi = upper ? upper[c] : INFINITY;
              yr[c] = clipf(mean_fn[c] + ax[j], lo, hi);
              if (dfeat) {
                const float lo2 = lower ? lower[d + c] : -INFINITY;
                const float hi2 = upper ? upper[d + c] : INFINITY;
                yr[d + c] = clipf(ad[j], lo2, hi2);
              }
            }
          }
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// Feature sampler, register-tiled (d <= 4): one thread per time step of a 256-step tile, 32
// samples per block iteration.  The tiny w = mean + L_in z L_out^T of the 32 samples goes to
// shared memory as [f][sample][c]; the output product then reads, per feature, one Phi / Phi'
// value and the w of 8 samples (broadcast 16-byte loads) for 8 * dy FMAs, and every thread stores
// its dy outputs of one (sample, time step) as one vector -- a warp writes 32 * dy * 4 contiguous
// bytes.  HBM-bound on the (K, T, dy) write when F is small (C4: F = 21, dy = 4).
// ------------------------------------------------------------------------------------------
constexpr int kFeatTile = 256;  // time steps per block (= threads)
constexpr int kFeatS = 32;      // samples per block iteration
constexpr int kFeatSB = 8;      // samples per register block

__host__ __device__ inline int pad4(int n) { return (n + 3) & ~3; }

template <int DY>
__device__ __forceinline__ void store_row(float* dst, const float (&v)[DY]) {
  if constexpr (DY % 4 == 0) {
#pragma unroll
    for (int i = 0; i < DY; i += 4)
      *reinterpret_cast<float4*>(dst + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
  } else if constexpr (DY % 2 == 0) {
#pragma unroll
    for (int i = 0; i < DY; i += 2) *reinterpret_cast<float2*>(dst + i) = make_float2(v[i], v[i + 1]);
  } else {
#pragma unroll
    for (int i = 0; i < DY; ++i) dst[i] = v[i];
  }
}

template <int DD, bool DERIV>
__global__ void __launch_bounds__(kFeatTile)
sample_features_tiled_kernel(const float* __restrict__ z, uint64_t seed, uint64_t draw,
                             uint64_t row_offset, int64_t K, int F, int T,
                             const float* __restrict__ mean_fn, const float* __restrict__ mean,
                             const float* __restrict__ l_in, const float* __restrict__ l_out,
                             const float* __restrict__ feat, const float* __restrict__ dfeat,
                             const float* __restrict__ lower, const float* __restrict__ upper,
                             float* __restrict__ w, float* __restrict__ y) {
  extern __shared__ __align__(16) float smem[];
  constexpr int DY = DERIV ? 2 * DD : DD;
  const int FD = F * DD;
  float* sPhi = smem;                                   // F * 256 (transposed tile)
  float* sdPhi = sPhi + F * kFeatTile;                  // F * 256 (DERIV)
  float* sLin = sdPhi + (DERIV ? F * kFeatTile : 0);    // F * F
  float* sLout = sLin + pad4(F * F);                    // d * d
  float* sMean = sLout + pad4(DD * DD);                 // F * d
  float* sZ = sMean + pad4(FD);                         // 32 * F * d
  float* sU = sZ + pad4(kFeatS * FD);                   // 32 * F * d
  float* sW = sU + pad4(kFeatS * FD);                   // [F][32][d]

  const int t0 = blockIdx.y * kFeatTile;
  const int nt = min(kFeatTile, T - t0);
  for (int i = threadIdx.x; i < F * kFeatTile; i += blockDim.x) {
    const int f = i / kFeatTile, tl = i % kFeatTile;
    const bool ok = tl < nt;
    sPhi[i] = ok ? feat[(int64_t)(t0 + tl) * F + f] : 0.f;
    if (DERIV) sdPhi[i] = ok ? dfeat[(int64_t)(t0 + tl) * F + f] : 0.f;
  }
  for (int i = threadIdx.x; i < F * F; i += blockDim.x) sLin[i] = l_in[i];
  for (int i = threadIdx.x; i < DD * DD; i += blockDim.x) sLout[i] = l_out[i];
  for (int i = threadIdx.x; i < FD; i += blockDim.x) sMean[i] = mean[i];
  float lo[DY], hi[DY], mf[DY];
#pragma unroll
  for (int c = 0; c < DY; ++c) {
    lo[c] = lower ? lower[c] : -INFINITY;
    hi[c] = upper ? upper[c] : INFINITY;
    mf[c] = c < DD ? mean_fn[c] : 0.f;
  }
  __syncthreads();

  const int tl = threadIdx.x;
  const bool active = tl < nt;
  const int64_t n_groups = ceil_div(K, kFeatS);
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t b0 = g * kFeatS;
    const int ns = (int)min((int64_t)kFeatS, K - b0);
    load_normals(sZ, ns * FD, z, (uint64_t)b0 * FD, (row_offset + (uint64_t)b0) * FD, seed, draw);
    __syncthreads();
    for (int e = threadIdx.x; e < ns * FD; e += blockDim.x) {
      const int l = e % DD;
      const float* zr = sZ + (e - l);
      float u = 0.f;
#pragma unroll
      for (int j = 0; j < DD; ++j) u = fmaf(zr[j], sLout[l * DD + j], u);
      sU[e] = u;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < kFeatS * FD; e += blockDim.x) {
      const int s = e / FD, r = e % FD, k = r / DD, l = r % DD;
      float a = 0.f;
      if (s < ns) {
        const float* row = sLin + k * F;
        for (int i = 0; i <= k; ++i) a = fmaf(row[i], sU[s * FD + i * DD + l], a);
        a += sMean[r];
        if (blockIdx.y == 0) w[(b0 + s) * FD + r] = a;
      }
      sW[(k * kFeatS + s) * DD + l] = a;
    }
    __syncthreads();
    for (int sb = 0; sb * kFeatSB < ns; ++sb) {
      float acc[kFeatSB][DY];
#pragma unroll
      for (int s = 0; s < kFeatSB; ++s)
#pragma unroll
        for (int c = 0; c < DY; ++c) acc[s][c] = 0.f;
      for (int f = 0; f < F; ++f) {
        const float ph = sPhi[f * kFeatTile + tl];
        const float dph = DERIV ? sdPhi[f * kFeatTile + tl] : 0.f;
        float wv[kFeatSB * DD];
        const float4* wp = reinterpret_cast<const float4*>(sW + (f * kFeatS + sb * kFeatSB) * DD);
#pragma unroll
        for (int q = 0; q < kFeatSB * DD / 4; ++q) {
          const float4 t = wp[q];
          wv[4 * q] = t.x; wv[4 * q + 1] = t.y; wv[4 * q + 2] = t.z; wv[4 * q + 3] = t.w;
        }
#pragma unroll
        for (int s = 0; s < kFeatSB; ++s)
#pragma unroll
          for (int c = 0; c < DD; ++c) {
            acc[s][c] = fmaf(ph, wv[s * DD + c], acc[s][c]);
            if (DERIV) acc[s][DD + c] = fmaf(dph, wv[s * DD + c], acc[s][DD + c]);
          }
      }
      if (active) {
#pragma unroll
        for (int s = 0; s < kFeatSB; ++s) {
          const int sl = sb * kFeatSB + s;
          if (sl < ns) {
            float out[DY];
#pragma unroll
            for (int c = 0; c < DY; ++c) out[c] = clipf(mf[c] + acc[s][c], lo[c], hi[c]);
            store_row<DY>(y + ((b0 + sl) * (int64_t)T + (t0 + tl)) * DY, out);
          }
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// Weights of a feature policy alone: w = mean + L_in z L_out^T, (K, F, d)  (ppi/policies.py:1009).
// First half of the two-kernel feature sampler: the expansion y = [Phi w, Phi' w] then runs on the
// tensor cores (features_expand_mma_kernel, mma_small.cu) reading w back (1 % of its traffic).
// Same arithmetic, Philox element order and summation order as the fused kernels above.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
feature_weights_kernel(const float* __restrict__ z, uint64_t seed, uint64_t draw, uint64_t row_offset,
                       int64_t K, int F, int d, int S, const float* __restrict__ mean,
                       const float* __restrict__ l_in, const float* __restrict__ l_out,
                       float* __restrict__ w) {
  extern __shared__ __align__(16) float smem[];
  const int FD = F * d;
  float* sLin = smem;                    // F * F
  float* sLout = sLin + pad4(F * F);     // d * d
  float* sMean = sLout + pad4(d * d);    // F * d
  float* sZ = sMean + pad4(FD);          // S * F * d
  float* sU = sZ + pad4(S * FD);         // S * F * d
  for (int i = threadIdx.x; i < F * F; i += blockDim.x) sLin[i] = l_in[i];
  for (int i = threadIdx.x; i < d * d; i += blockDim.x) sLout[i] = l_out[i];
  for (int i = threadIdx.x; i < FD; i += blockDim.x) sMean[i] = mean[i];
  // static work split: thread = (block of 8 samples, row pair (p, F-1-p), output column l).  L_in is
  // lower triangular, so rows p and F-1-p together cost F + 1 multiply-adds for every thread (no
  // divergence on the triangle), and each L_in element read from shared memory feeds 8 samples.
  // S = 8 * (256 / (ceil(F/2) d)) samples per group.
  const int P = (F + 1) >> 1, PD = P * d;
  const int nsub = S >> 3;
  const int sub = threadIdx.x / PD, rem = threadIdx.x - sub * PD, p = rem / d, l = rem - p * d;
  const int k1 = p, k2 = F - 1 - p;
  const bool worker = sub < nsub;
  const int64_t n_groups = ceil_div(K, S);
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t b0 = g * S;
    const int ns = (int)min((int64_t)S, K - b0);
    __syncthreads();
    load_normals(sZ, ns * FD, z, (uint64_t)b0 * FD, (row_offset + (uint64_t)b0) * FD, seed, draw);
    __syncthreads();
    for (int e = threadIdx.x; e < ns * FD; e += blockDim.x) {
      const int c = e % d;
      const float* zr = sZ + (e - c);
      float u = 0.f;
      for (int j = 0; j < d; ++j) u = fmaf(zr[j], sLout[c * d + j], u);
      sU[e] = u;
    }
    __syncthreads();
    if (worker && sub * 8 < ns) {
      float a1[8], a2[8];
      const float* up[8];
#pragma unroll
      for (int m = 0; m < 8; ++m) {
        a1[m] = a2[m] = 0.f;
        up[m] = sU + min(sub * 8 + m, ns - 1) * FD + l;
      }
      const float* row1 = sLin + k1 * F;
      const float* row2 = sLin + k2 * F;
      for (int i = 0; i <= k2; ++i) {
        const float li2 = row2[i];
        const float li1 = i <= k1 ? row1[i] : 0.f;   // a zero coefficient leaves the sum bit-identical
#pragma unroll
        for (int m = 0; m < 8; ++m) {
          const float u = up[m][i * d];
          a2[m] = fmaf(li2, u, a2[m]);
          a1[m] = fmaf(li1, u, a1[m]);
        }
      }
      const int r1 = k1 * d + l, r2 = k2 * d + l;
      const float m1 = sMean[r1], m2 = sMean[r2];
#pragma unroll
      for (int m = 0; m < 8; ++m) {
        const int sm = sub * 8 + m;
        if (sm < ns) {
          float* wr = w + (b0 + sm) * FD;
          wr[r2] = a2[m] + m2;
          if (k1 != k2) wr[r1] = a1[m] + m1;
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// iid sampler: element-wise, one quad of the base-normal matrix per thread iteration.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sample_iid_kernel(const float* __restrict__ z, uint64_t seed, uint64_t draw, uint64_t row_offset,
                  int64_t K, int HD, int d, const float* __restrict__ mean_fn,
                  const float* __restrict__ mean, const float* __restrict__ stdv,
                  const float* __restrict__ lower, const float* __restrict__ upper,
                  float* __restrict__ x) {
  const int64_t n_local = K * HD;
  const uint64_t elem_base = row_offset * (uint64_t)HD;
  const uint64_t q0 = elem_base >> 2, q1 = (elem_base + n_local + 3) >> 2;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t q = q0 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < q1; q += stride) {
    float v[4];
    const int64_t local = (int64_t)(q * 4) - (int64_t)elem_base;
    if (z == nullptr) philox_normal4(seed, draw, q, v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t l = local + i;
      if (l >= 0 && l < n_local) {
        const float zz = z ? z[l] : v[i];
        const int e = (int)(l % HD), c = e % d;
        const float lo = lower ? lower[c] : -INFINITY, hi = upper ? upper[c] : INFINITY;
        x[l] = clipf(mean_fn[c] + mean[e] + stdv[e] * zz, lo, hi);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// Small dense Gaussian transform x = mu + z L^T, D <= 128: L^T in shared memory.
// ------------------------------------------------------------------------------------------
template <int S>
__global__ void __launch_bounds__(256)
sample_gaussian_small_kernel(const float* __restrict__ z, uint64_t seed, uint64_t draw,
                             uint64_t row_offset, int64_t K, int D, const float* __restrict__ mu,
                             const float* __restrict__ L, float* __restrict__ x) {
  extern __shared__ float smem[];
  float* sLt = smem;            // D*D, sLt[j*D+i] = L[i][j]
  float* sZ = sLt + D * D;      // S*D
  for (int e = threadIdx.x; e < D * D; e += blockDim.x) {
    const int i = e / D, j = e % D;
    sLt[j * D + i] = L[e];
  }
  __syncthreads();
  const int64_t n_groups = ceil_div(K, S);
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t b0 = g * S;
    const int ns = (int)min((int64_t)S, K - b0);
    load_normals(sZ, ns * D, z, (uint64_t)b0 * D, (row_offset + (uint64_t)b0) * D, seed, draw);
    __syncthreads();
    for (int e = threadIdx.x; e < ns * D; e += blockDim.x) {
      const int s = e / D, i = e % D;
      float a = 0.f;
      const float* zr = sZ + s * D;
      for (int j = 0; j <= i; ++j) a = fmaf(zr[j], sLt[j * D + i], a);
      x[(b0 + s) * D + i] = mu[i] + a;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// Scrambled-Sobol base normals (QuasiMonteCarlo, ppi/samplers.py:25-44): out[r][c] =
// sqrt(2) erfinv(2 shrinkage x_{first + r, c} - 1) with x_i = 2^-bits (shift ^ XOR_{b in gray(i)} sv[c][b]),
// the closed form of the Gray-code recurrence of scipy.stats.qmc.Sobol (direction numbers with the
// Owen-type scrambling already folded into sv / shift by the engine's constructor).  A block stages
// the direction numbers of 128 columns in shared memory ([bit][column]: conflict-free); each thread
// owns one column, builds its first point from the set bits of the Gray code and then walks down
// its rows with one XOR per point (the bit that flips between i and i + 1 is the lowest zero bit of
// i).  The inverse error function runs in fp64: 2 x - 1 needs all 30 bits near the tails.
// ------------------------------------------------------------------------------------------
constexpr int kSobolCols = 128;   // columns per block (= threads)
constexpr int kSobolRows = 64;    // consecutive rows per thread

__global__ void __launch_bounds__(kSobolCols)
sobol_normal_kernel(float* __restrict__ out, int64_t rows, int64_t cols, const uint32_t* __restrict__ sv,
                    const uint32_t* __restrict__ shift, int bits, uint64_t first, double shrink_scale) {
  __shared__ uint32_t s_sv[32 * kSobolCols];
  const int64_t c0 = (int64_t)blockIdx.x * kSobolCols;
  const int nc = (int)min((int64_t)kSobolCols, cols - c0);
  for (int e = threadIdx.x; e < nc * bits; e += blockDim.x) {
    const int c = e / bits, b = e - c * bits;
    s_sv[b * kSobolCols + c] = sv[(c0 + c) * bits + b];
  }
  __syncthreads();
  if ((int)threadIdx.x >= nc) return;
  const int64_t c = c0 + threadIdx.x;
  const uint32_t sh = shift[c];
  for (int64_t r0 = (int64_t)blockIdx.y * kSobolRows; r0 < rows; r0 += (int64_t)gridDim.y * kSobolRows) {
    uint64_t i = first + (uint64_t)r0;
    uint64_t gray = i ^ (i >> 1);
    uint32_t q = sh;
    while (gray) {
      const int b = __ffsll((long long)gray) - 1;
      gray &= gray - 1;
      q ^= s_sv[b * kSobolCols + threadIdx.x];
    }
    const int nr = (int)min((int64_t)kSobolRows, rows - r0);
    float* o = out + r0 * cols + c;
    for (int r = 0; r < nr; ++r) {
      const double u = (double)q * shrink_scale;                      // shrinkage * x
      o[(int64_t)r * cols] = (float)(1.4142135623730951 * erfinv(2.0 * u - 1.0));
      const int b = __ffsll((long long)~i) - 1;                       // lowest zero bit of i
      q ^= s_sv[min(b, 31) * kSobolCols + threadIdx.x];
      ++i;
    }
  }
}

// Cubature sigma points sqrt(d) [I; -I]  (CubatureQuadrature, ppi/samplers.py:47-53), (2 d, d).
__global__ void __launch_bounds__(256)
cubature_points_kernel(float* __restrict__ out, int d) {
  const float s = sqrtf((float)d);
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < 2 * (int64_t)d * d;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / d, c = e - r * d;
    out[e] = (r == c) ? s : (r - d == c ? -s : 0.f);
  }
}

// ------------------------------------------------------------------------------------------
// Kernel Gram matrix + Cholesky, fp64, one block.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double gram_entry(int kind, double a, double b, double sigma, double ls,
                                             double period, bool square, bool on_diag) {
  const double diff = a - b;
  switch (kind) {
    case PPI_KERNEL_SE: {
      const double r = diff / ls;
      double k = sigma * exp(-0.5 * r * r);
      if (square && on_diag) k += 1e-3 * sigma;  // ppi/policies.py:929-931
      return k;
    }
    case PPI_KERNEL_MATERN12: {
      double ad = fabs(diff);
      if (ad == 0.0) ad = 1e-8;  // ppi/policies.py:941
      return sigma * exp(-ad / ls);
    }
    case PPI_KERNEL_MATERN32: {
      double ad = fabs(diff);
      if (ad == 0.0) ad = 1e-8;
      const double dd = sqrt(3.0) * ad / ls;
      return sigma * (1.0 + dd) * exp(-dd);
    }
    case PPI_KERNEL_MATERN52: {
      double ad = fabs(diff);
      if (ad == 0.0) ad = 1e-8;
      const double dd = sqrt(5.0) * ad / ls;
      return sigma * (1.0 + dd + dd * dd / 3.0) * exp(-dd);
    }
    case PPI_KERNEL_PERIODIC: {
      double ad = fabs(diff == 0.0 ? -1e-8 : diff);  // ppi/policies.py:1031-1032
      const double sn = sin(M_PI * ad / period);
      double k = sigma * exp(-2.0 * sn * sn / ls);
      if (square && on_diag) k += 1e-3 * sigma;
      return k;
    }
    default:  // white noise, ppi/policies.py:1087-1091
      return diff == 0.0 ? sigma : 0.0;
  }
}

__global__ void __launch_bounds__(256)
gram_chol_kernel(const double* __restrict__ t1, int n1, const double* __restrict__ t2, int n2,
                 int kind, double sigma, double ls, double period, double* __restrict__ gram64,
                 float* __restrict__ gram32, double* __restrict__ chol64,
                 float* __restrict__ chol32, int* __restrict__ status) {
  extern __shared__ double sa[];
  __shared__ int fail;
  const bool square = (n1 == n2);
  const bool do_chol = (chol64 != nullptr || chol32 != nullptr);
  for (int e = threadIdx.x; e < n1 * n2; e += blockDim.x) {
    const int i = e / n2, j = e % n2;
    const double k = gram_entry(kind, t1[i], t2[j], sigma, ls, period, square, i == j);
    if (gram64) gram64[e] = k;
    if (gram32) gram32[e] = (float)k;
    if (do_chol) sa[e] = k;
  }
  if (!do_chol) return;
  __syncthreads();
  block_cholesky(sa, n1, &fail);
  for (int e = threadIdx.x; e < n1 * n1; e += blockDim.x) {
    const int i = e / n1, j = e % n1;
    const double v = (j <= i) ? sa[e] : 0.0;
    if (chol64) chol64[e] = v;
    if (chol32) chol32[e] = (float)v;
  }
  if (threadIdx.x == 0 && status) *status = fail;
}

// ------------------------------------------------------------------------------------------
// MPC time shift of a kernel prior (BaseKernel.update_timesteps, ppi/policies.py:761-791), fp64,
// one block: the step mpc/mpc.py:64-65 runs before every planner update.
//   gain  = P^-1 (P - Sigma) P^-T                 covariance reduction of the old window
//   cross = k(t_new, t_old)
//   mean' = clip(cross P^-1 mean + mean_fn) - mean_fn
//   cov'  = k(t_new, t_new) - anneal cross gain cross^T + jitter I,   chol' = cholesky(cov')
// Five H x H products, the two Gram matrices and the factorisation in ONE launch (the host mirror
// needed three Gram launches with a device->host read each plus NumPy).  The matrices live in a
// global workspace (L1/L2 resident at these sizes); cov' is factorised in shared memory.
// ------------------------------------------------------------------------------------------
// C (m x n) = A (m x k) B (k x n);  TB: B is given transposed, (n x k).  A single SM moves 128 B
// per clock through L1, i.e. 16 B per thread and clock at most: a one-output-per-thread product
// (16 B of loads per FMA) is bound by that, not by the fp64 pipe.  Each thread therefore owns a
// 4 x 4 block of C (8 loads per 16 FMAs).  (No __restrict__ / read-only loads: the operands may
// have been written earlier in this kernel.)
template <bool TB>
__device__ __forceinline__ void block_gemm64(const double* A, const double* B, double* C, int m, int n,
                                             int k) {
  const int tm = (m + 3) >> 2, tn = (n + 3) >> 2;
  for (int e = threadIdx.x; e < tm * tn; e += blockDim.x) {
    // interleaved ownership: rows ti + tm r, columns tj + tn c, so that the threads of a warp read
    // consecutive columns of B (one or two 128-byte lines per load) and mostly the same rows of A
    const int ti = e / tn, tj = e - ti * tn;
    const double* a[4];
    const double* b[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      a[r] = A + (size_t)min(ti + tm * r, m - 1) * k;
      const int c = min(tj + tn * r, n - 1);
      b[r] = TB ? B + (size_t)c * k : B + c;
    }
    const size_t sb = TB ? 1 : (size_t)n;
    double acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[r][c] = 0.0;
    for (int l = 0; l < k; ++l) {
      double av[4], bv[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        av[r] = a[r][l];
        bv[r] = b[r][(size_t)l * sb];
      }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = fma(av[r], bv[c], acc[r][c]);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (ti + tm * r < m && tj + tn * c < n) C[(size_t)(ti + tm * r) * n + tj + tn * c] = acc[r][c];
  }
  __syncthreads();
}

__global__ void __launch_bounds__(512)
kernel_shift_kernel(int kind, double sigma, double ls, double period, const double* __restrict__ t_old,
                    int n_old, const double* __restrict__ t_new, int n_new,
                    const double* __restrict__ prior, const double* __restrict__ prior_inv,
                    const double* __restrict__ cov_in, const double* __restrict__ mean, int d,
                    const double* __restrict__ mean_fn, const double* __restrict__ lower,
                    const double* __restrict__ upper, double anneal, double jitter,
                    double* __restrict__ mean_new, double* __restrict__ cov_new,
                    double* __restrict__ chol_new, double* __restrict__ status, double* ws) {
  extern __shared__ double sa[];   // n_new x n_new
  __shared__ int fail;
  double* w0 = ws;                                  // n_old^2: P - Sigma, later gain
  double* w1 = w0 + (size_t)n_old * n_old;          // n_old^2: P^-1 (P - Sigma)
  double* cross = w1 + (size_t)n_old * n_old;       // n_new x n_old
  double* w2 = cross + (size_t)n_new * n_old;       // n_new x n_old: cross P^-1, later cross gain
  for (int e = threadIdx.x; e < n_old * n_old; e += blockDim.x) w0[e] = prior[e] - cov_in[e];
  for (int e = threadIdx.x; e < n_new * n_old; e += blockDim.x) {
    const int i = e / n_old, j = e - i * n_old;
    cross[e] = gram_entry(kind, t_new[i], t_old[j], sigma, ls, period, n_new == n_old, i == j);
  }
  __syncthreads();
  block_gemm64<false>(prior_inv, w0, w1, n_old, n_old, n_old);
  block_gemm64<true>(w1, prior_inv, w0, n_old, n_old, n_old);       // gain = w1 P^-T
  block_gemm64<false>(cross, prior_inv, w2, n_new, n_old, n_old);   // cross P^-1
  for (int e = threadIdx.x; e < n_new * d; e += blockDim.x) {
    const int i = e / d, c = e - i * d;
    double acc = 0.0;
    for (int l = 0; l < n_old; ++l) acc = fma(w2[(size_t)i * n_old + l], mean[(size_t)l * d + c], acc);
    double v = acc + mean_fn[c];
    if (lower) v = fmin(fmax(v, lower[c]), upper[c]);
    mean_new[e] = v - mean_fn[c];
  }
  __syncthreads();
  block_gemm64<false>(cross, w0, w2, n_new, n_old, n_old);          // cross gain
  block_gemm64<true>(w2, cross, cov_new, n_new, n_new, n_old);      // (cross gain) cross^T
  for (int e = threadIdx.x; e < n_new * n_new; e += blockDim.x) {
    const int i = e / n_new, j = e - i * n_new;
    double v = gram_entry(kind, t_new[i], t_new[j], sigma, ls, period, true, i == j) - anneal * cov_new[e];
    if (i == j) v += jitter;
    cov_new[e] = v;
    sa[e] = v;
  }
  __syncthreads();
  block_cholesky(sa, n_new, &fail);
  for (int e = threadIdx.x; e < n_new * n_new; e += blockDim.x) {
    const int i = e / n_new, j = e - i * n_new;
    chol_new[e] = (j <= i) ? sa[e] : 0.0;
  }
  if (threadIdx.x == 0) *status = (double)fail;
}

__global__ void __launch_bounds__(256)
chol_small_kernel(const double* __restrict__ a, int n, double* __restrict__ l64,
                  float* __restrict__ l32, int* __restrict__ status) {
  extern __shared__ double sa[];
  __shared__ int fail;
  for (int e = threadIdx.x; e < n * n; e += blockDim.x) sa[e] = a[e];
  __syncthreads();
  block_cholesky(sa, n, &fail);
  for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
    const int i = e / n, j = e % n;
    const double v = (j <= i) ? sa[e] : 0.0;
    if (l64) l64[e] = v;
    if (l32) l32[e] = (float)v;
  }
  if (threadIdx.x == 0 && status) *status = fail;
}

constexpr int kMaxSmemBytes = 200 * 1024;

template <typename Kern>
static int set_smem(Kern kern, size_t bytes) {
  if (bytes > 48 * 1024) {
    PPI_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  }
  return 0;
}

static int grid_for(int64_t work_items, int per_sm) {
  int64_t g = work_items < 1 ? 1 : work_items;
  const int64_t cap = (int64_t)kNumSMs * per_sm;
  return (int)(g < cap ? g : cap);
}

}  // namespace ppi

using namespace ppi;

// mma_small.cu
int ppi_sample_matnorm_mma(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset, int64_t K,
                           int H, int d, const float* mean_fn, const float* mean, const float* l_in,
                           const float* l_out, const float* lower, const float* upper, float* x,
                           cudaStream_t st);
int ppi_features_expand_mma(const float* w, int64_t K, int F, int T, const float* mean_fn,
                            const float* feat, const float* dfeat, const float* lower,
                            const float* upper, float* y, cudaStream_t st);

extern "C" int ppi_philox_normal(float* out, int64_t rows, int64_t cols, uint64_t seed,
                                 uint64_t draw_id, uint64_t row_offset, void* stream) {
  PPI_REQUIRE(out && rows >= 0 && cols > 0, PPI_ERR_INVALID, "ppi_philox_normal: bad arguments");
  if (rows == 0) return 0;
  const int64_t n = rows * cols;
  const int grid = grid_for(ceil_div(ceil_div(n, 4), 256 * 4), 8);
  philox_normal_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(out, n, row_offset * (uint64_t)cols,
                                                                      seed, draw_id);
  return check_launch("philox_normal_kernel");
}

extern "C" int ppi_philox_normal_full(float* out, int64_t rows, int64_t cols, uint64_t seed,
                                      uint64_t draw_id, uint64_t row_offset, void* stream) {
  PPI_REQUIRE(out && rows >= 0 && cols > 0, PPI_ERR_INVALID, "ppi_philox_normal_full: bad arguments");
  if (rows == 0) return 0;
  const int64_t n = rows * cols;
  const int grid = grid_for(ceil_div(ceil_div(n, 4), 256 * 4), 8);
  philox_normal_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(out, n, row_offset * (uint64_t)cols,
                                                                     seed, draw_id);
  return check_launch("philox_normal_kernel<full>");
}

extern "C" int ppi_kernel_gram_chol(const double* t1, int n1, const double* t2, int n2, int kind,
                                    double sigma, double lengthscale, double period, double* gram64,
                                    float* gram32, double* chol64, float* chol32, int* status,
                                    void* stream) {
  PPI_REQUIRE(t1 && t2 && n1 > 0 && n2 > 0, PPI_ERR_INVALID, "ppi_kernel_gram_chol: bad arguments");
  PPI_REQUIRE(kind >= 0 && kind <= PPI_KERNEL_WHITE, PPI_ERR_INVALID, "unknown kernel kind %d", kind);
  const bool do_chol = chol64 || chol32;
  size_t smem = 0;
  if (do_chol) {
    PPI_REQUIRE(n1 == n2, PPI_ERR_INVALID, "Cholesky needs a square Gram matrix");
    smem = (size_t)n1 * n1 * sizeof(double);
    PPI_REQUIRE(smem <= (size_t)kMaxSmemBytes, PPI_ERR_UNSUPPORTED,
                "single-block Cholesky supports n <= 160 (got %d)", n1);
    if (int rc = set_smem(gram_chol_kernel, smem)) return rc;
  }
  gram_chol_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(t1, n1, t2, n2, kind, sigma, lengthscale,
                                                           period, gram64, gram32, chol64, chol32,
                                                           status);
  return check_launch("gram_chol_kernel");
}

extern "C" int ppi_sobol_normal(float* out, int64_t rows, int64_t cols, const uint32_t* sv,
                                const uint32_t* shift, int bits, uint64_t first_index, double shrinkage,
                                void* stream) {
  PPI_REQUIRE(out && sv && shift && rows >= 0 && cols > 0, PPI_ERR_INVALID, "ppi_sobol_normal: bad arguments");
  PPI_REQUIRE(bits > 0 && bits <= 32, PPI_ERR_INVALID, "ppi_sobol_normal: bits must lie in [1, 32] (got %d)", bits);
  PPI_REQUIRE(first_index + (uint64_t)rows <= (1ull << bits), PPI_ERR_INVALID,
              "ppi_sobol_normal: the sequence holds 2^%d points", bits);
  if (rows == 0) return 0;
  const int64_t gx = ceil_div(cols, kSobolCols);
  int64_t gy = ceil_div(rows, kSobolRows);
  const int64_t cap = ceil_div((int64_t)kNumSMs * 16, gx);
  if (gy > cap) gy = cap;
  if (gy > 65535) gy = 65535;
  dim3 grid((unsigned)gx, (unsigned)gy);
  sobol_normal_kernel<<<grid, kSobolCols, 0, (cudaStream_t)stream>>>(out, rows, cols, sv, shift, bits, first_index,
                                                                     shrinkage * ldexp(1.0, -bits));
  return check_launch("sobol_normal_kernel");
}

extern "C" int ppi_cubature_points(float* out, int d, void* stream) {
  PPI_REQUIRE(out && d > 0, PPI_ERR_INVALID, "ppi_cubature_points: bad arguments");
  const int grid = (int)min((int64_t)kNumSMs * 4, ceil_div(2 * (int64_t)d * d, 256));
  cubature_points_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(out, d);
  return check_launch("cubature_points_kernel");
}

extern "C" size_t ppi_kernel_shift_workspace_bytes(int n_old, int n_new) {
  if (n_old <= 0 || n_new <= 0) return 0;
  return (2 * (size_t)n_old * n_old + 2 * (size_t)n_new * n_old) * sizeof(double);
}

extern "C" int ppi_kernel_shift(int kind, double sigma, double lengthscale, double period,
                                const double* t_old, int n_old, const double* t_new, int n_new,
                                const double* prior, const double* prior_inv, const double* cov_in,
                                const double* mean, int d, const double* mean_fn, const double* lower,
                                const double* upper, double anneal, double jitter, double* mean_new,
                                double* cov_new, double* chol_new, double* status, void* workspace,
                                void* stream) {
  PPI_REQUIRE(t_old && t_new && n_old > 0 && n_new > 0 && d > 0 && prior && prior_inv && cov_in && mean &&
                  mean_fn && mean_new && cov_new && chol_new && status && workspace,
              PPI_ERR_INVALID, "ppi_kernel_shift: bad arguments");
  PPI_REQUIRE(kind >= 0 && kind <= PPI_KERNEL_WHITE, PPI_ERR_INVALID, "unknown kernel kind %d", kind);
  PPI_REQUIRE((lower == nullptr) == (upper == nullptr), PPI_ERR_INVALID,
              "lower and upper must both be given or both be NULL");
  PPI_REQUIRE(anneal >= 0.0 && anneal <= 1.0, PPI_ERR_INVALID, "anneal must lie in [0, 1]");
  const size_t smem = (size_t)n_new * n_new * sizeof(double);
  PPI_REQUIRE(smem <= (size_t)kMaxSmemBytes, PPI_ERR_UNSUPPORTED,
              "single-block Cholesky supports n <= 160 (got %d)", n_new);
  if (int rc = set_smem(kernel_shift_kernel, smem)) return rc;
  kernel_shift_kernel<<<1, 512, smem, (cudaStream_t)stream>>>(
      kind, sigma, lengthscale, period, t_old, n_old, t_new, n_new, prior, prior_inv, cov_in, mean, d,
      mean_fn, lower, upper, anneal, jitter, mean_new, cov_new, chol_new, status, (double*)workspace);
  return check_launch("kernel_shift_kernel");
}

extern "C" int ppi_cholesky_small(const double* a, int n, double* l64, float* l32, int* status,
                                  void* stream) {
  PPI_REQUIRE(a && n > 0, PPI_ERR_INVALID, "ppi_cholesky_small: bad arguments");
  const size_t smem = (size_t)n * n * sizeof(double);
  PPI_REQUIRE(smem <= (size_t)kMaxSmemBytes, PPI_ERR_UNSUPPORTED,
              "single-block Cholesky supports n <= 160 (got %d)", n);
  if (int rc = set_smem(chol_small_kernel, smem)) return rc;
  chol_small_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(a, n, l64, l32, status);
  return check_launch("chol_small_kernel");
}

extern "C" int ppi_sample_matnorm(const float* z, uint64_t seed, uint64_t draw_id,
                                  uint64_t row_offset, int64_t K, int H, int d,
                                  const float* mean_fn, const float* mean, const float* l_in,
                                  const float* l_out, int l_out_is_diag, const float* lower,
                                  const float* upper, float* x, void* stream) {
  PPI_REQUIRE(K >= 0 && H > 0 && d > 0 && mean_fn && mean && l_in && l_out && x, PPI_ERR_INVALID,
              "ppi_sample_matnorm: bad arguments");
  PPI_REQUIRE((lower == nullptr) == (upper == nullptr), PPI_ERR_INVALID,
              "lower and upper must both be given or both be NULL");
  if (K == 0) return 0;
  if (H <= 32 && d <= 32 && l_out_is_diag) {  // warp-level tensor-core kernel (mma_small.cu)
    ProfileScope prof(PPI_PROF_SAMPLE_SMALL, (cudaStream_t)stream);
    return ppi_sample_matnorm_mma(z, seed, draw_id, row_offset, K, H, d, mean_fn, mean, l_in, l_out, lower,
                                  upper, x, (cudaStream_t)stream);
  }
  const size_t fixed = ((size_t)H * H + (size_t)d * d + (size_t)H * d) * sizeof(float);
  auto need = [&](int S) { return fixed + 2 * (size_t)S * H * d * sizeof(float); };
  cudaStream_t st = (cudaStream_t)stream;
#define PPI_LAUNCH_MATNORM(S)                                                                   \
  do {                                                                                          \
    if (int rc = set_smem(sample_matnorm_kernel<S>, need(S))) return rc;                        \
    const int grid = grid_for(ceil_div(K, S), 4);                                               \
    sample_matnorm_kernel<S><<<grid, 256, need(S), st>>>(z, seed, draw_id, row_offset, K, H, d, \
                                                         mean_fn, mean, l_in, l_out,            \
                                                         l_out_is_diag, lower, upper, x);       \
  } while (0)
  if (need(4) <= 48 * 1024) PPI_LAUNCH_MATNORM(4);
  else if (need(2) <= 100 * 1024) PPI_LAUNCH_MATNORM(2);
  else {
    PPI_REQUIRE(need(1) <= (size_t)kMaxSmemBytes, PPI_ERR_UNSUPPORTED,
                "ppi_sample_matnorm: H=%d d=%d does not fit shared memory", H, d);
    PPI_LAUNCH_MATNORM(1);
  }
#undef PPI_LAUNCH_MATNORM
  return check_launch("sample_matnorm_kernel");
}

extern "C" int ppi_sample_features(const float* z, uint64_t seed, uint64_t draw_id,
                                   uint64_t row_offset, int64_t K, int F, int d, int T,
                                   const float* mean_fn, const float* mean, const float* l_in,
                                   const float* l_out, const float* feat, const float* dfeat,
                                   const float* lower, const float* upper, float* w, float* y,
                                   void* stream) {
  PPI_REQUIRE(K >= 0 && F > 0 && d > 0 && T > 0 && mean_fn && mean && l_in && l_out && feat && w && y,
              PPI_ERR_INVALID, "ppi_sample_features: bad arguments");
  PPI_REQUIRE((lower == nullptr) == (upper == nullptr), PPI_ERR_INVALID,
              "lower and upper must both be given or both be NULL");
  if (K == 0) return 0;
  if (d == 2 && F <= 32 && (reinterpret_cast<uintptr_t>(y) & 15) == 0) {
    // two kernels: weights (SIMT, tiny), then the expansion on mma.sync (mma_small.cu)
    const int FD = F * d;
    const int S = 8 * (256 / (((F + 1) / 2) * d));   // samples per group (>= 64 for F <= 32, d = 2)
    const size_t smem_w = ((size_t)pad4(F * F) + pad4(d * d) + pad4(FD) + 2 * (size_t)pad4(S * FD)) *
                          sizeof(float);
    if (int rc = set_smem(feature_weights_kernel, smem_w)) return rc;
    ProfileScope prof(PPI_PROF_SAMPLE_SMALL, (cudaStream_t)stream);
    feature_weights_kernel<<<grid_for(ceil_div(K, S), 4), 256, smem_w, (cudaStream_t)stream>>>(
        z, seed, draw_id, row_offset, K, F, d, S, mean, l_in, l_out, w);
    if (int rc = check_launch("feature_weights_kernel")) return rc;
    return ppi_features_expand_mma(w, K, F, T, mean_fn, feat, dfeat, lower, upper, y,
                                   (cudaStream_t)stream);
  }
  if (d <= 4) {
    const int FD = F * d;
    const size_t smem_t = ((size_t)(dfeat ? 2 : 1) * F * kFeatTile + pad4(F * F) + pad4(d * d) + pad4(FD) +
                           3 * (size_t)pad4(kFeatS * FD)) * sizeof(float);
    const bool y_ok = (reinterpret_cast<uintptr_t>(y) & 15) == 0;
    if (smem_t <= 100 * 1024 && y_ok) {
      const int ty = (int)ceil_div(T, kFeatTile);
      int gx = (int)ceil_div(grid_for(ceil_div(K, kFeatS), 3), ty);
      if (gx < 1) gx = 1;
      dim3 grid(gx, ty);
      ProfileScope prof(PPI_PROF_SAMPLE_SMALL, (cudaStream_t)stream);
#define PPI_LAUNCH_FEAT(DD, DERIV)                                                                  \
  do {                                                                                              \
    if (int rc = set_smem(sample_features_tiled_kernel<DD, DERIV>, smem_t)) return rc;              \
    sample_features_tiled_kernel<DD, DERIV><<<grid, kFeatTile, smem_t, (cudaStream_t)stream>>>(     \
        z, seed, draw_id, row_offset, K, F, T, mean_fn, mean, l_in, l_out, feat, dfeat, lower,      \
        upper, w, y);                                                                               \
  } while (0)
      if (dfeat) {
        if (d == 1) PPI_LAUNCH_FEAT(1, true); else if (d == 2) PPI_LAUNCH_FEAT(2, true);
        else if (d == 3) PPI_LAUNCH_FEAT(3, true); else PPI_LAUNCH_FEAT(4, true);
      } else {
        if (d == 1) PPI_LAUNCH_FEAT(1, false); else if (d == 2) PPI_LAUNCH_FEAT(2, false);
        else if (d == 3) PPI_LAUNCH_FEAT(3, false); else PPI_LAUNCH_FEAT(4, false);
      }
#undef PPI_LAUNCH_FEAT
      return check_launch("sample_features_tiled_kernel");
    }
  }
  constexpr int S = 4;
  const size_t smem = (2 * (size_t)F * kFeatTT + (size_t)F * F + (size_t)d * d + (size_t)F * d +
                       3 * (size_t)S * F * d) * sizeof(float);
  PPI_REQUIRE(smem <= (size_t)kMaxSmemBytes, PPI_ERR_UNSUPPORTED,
              "ppi_sample_features: F=%d d=%d does not fit shared memory", F, d);
  if (int rc = set_smem(sample_features_kernel<S>, smem)) return rc;
  const int ty = (int)ceil_div(T, kFeatTT);
  int gx = grid_for(ceil_div(K, S), 4);
  gx = (int)ceil_div(gx, ty);
  if (gx < 1) gx = 1;
  dim3 grid(gx, ty);
  ProfileScope prof(PPI_PROF_SAMPLE_SMALL, (cudaStream_t)stream);
  sample_features_kernel<S><<<grid, 256, smem, (cudaStream_t)stream>>>(
      z, seed, draw_id, row_offset, K, F, d, T, mean_fn, mean, l_in, l_out, feat, dfeat, lower, upper,
      w, y);
  return check_launch("sample_features_kernel");
}

extern "C" int ppi_sample_iid(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                              int64_t K, int H, int d, const float* mean_fn, const float* mean,
                              const float* stdv, const float* lower, const float* upper, float* x,
                              void* stream) {
  PPI_REQUIRE(K >= 0 && H > 0 && d > 0 && mean_fn && mean && stdv && x, PPI_ERR_INVALID,
              "ppi_sample_iid: bad arguments");
  PPI_REQUIRE((lower == nullptr) == (upper == nullptr), PPI_ERR_INVALID,
              "lower and upper must both be given or both be NULL");
  if (K == 0) return 0;
  const int64_t n = K * H * d;
  const int grid = grid_for(ceil_div(ceil_div(n, 4), 256 * 4), 8);
  sample_iid_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(z, seed, draw_id, row_offset, K, H * d, d,
                                                            mean_fn, mean, stdv, lower, upper, x);
  return check_launch("sample_iid_kernel");
}

// Called from tc_gemm.cu's ppi_sample_gaussian for D <= 128.
int ppi_sample_gaussian_small(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                              int64_t K, int D, const float* mu, const float* L, float* x,
                              cudaStream_t st) {
  constexpr int S = 8;
  const size_t smem = ((size_t)D * D + (size_t)S * D) * sizeof(float);
  if (int rc = set_smem(sample_gaussian_small_kernel<S>, smem)) return rc;
  const int grid = grid_for(ceil_div(K, S), 2);
  ProfileScope prof(PPI_PROF_SAMPLE_SMALL, st);
  sample_gaussian_small_kernel<S><<<grid, 256, smem, st>>>(z, seed, draw_id, row_offset, K, D, mu, L, x);
  return check_launch("sample_gaussian_small_kernel");
}
