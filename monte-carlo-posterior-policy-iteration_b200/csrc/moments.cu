// Weighted moment matching, HBM-bound / small-output forms.
//   first + element-wise second moment     ppi/policies.py:88,122,1158-1162
//   matrix-normal input covariance         ppi/policies.py:129-132  (m_projection_mavn)
//   small dense covariance (d = 1)         ppi/policies.py:89-92    (m_projection)
//   finalisers                             ppi/policies.py:53-62,92,132,480-481,819
// The large dense covariance (tcgen05) lives in tc_gemm.cu.
#include "common.cuh"

namespace ppi {

// ------------------------------------------------------------------------------------------
// s1[i] += sum_k nw_k (x_ki - a_i),  s2[i] += sum_k nw_k (x_ki - b_i)^2
// Block = CG column groups (VEC columns each) x RL row lanes; every thread streams its rows with
// VEC-wide loads, accumulates <= 64 rows in fp32, then folds into fp64; row lanes are combined
// through shared memory and one fp64 atomicAdd per column per block reaches global memory.
// ------------------------------------------------------------------------------------------
template <int VEC, bool SECOND>
__global__ void __launch_bounds__(256)
wmoments_diag_kernel(const float* __restrict__ x, const float* __restrict__ nw, int64_t K, int64_t D,
                     const float* __restrict__ shift1, const float* __restrict__ shift2,
                     double* __restrict__ s1, double* __restrict__ s2, unsigned int* __restrict__ absdev,
                     int cgb, int64_t rows_per_block) {
  __shared__ double red[256 * VEC];
  const int64_t ncg = D / VEC;
  const int cg_local = threadIdx.x % cgb, rl = threadIdx.x / cgb, nrl = blockDim.x / cgb;
  const int64_t cg = (int64_t)blockIdx.x * cgb + cg_local;
  const bool active = cg < ncg && rl < nrl;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_block;
  const int64_t r1 = min(K, r0 + rows_per_block);
  float a[VEC], b[VEC];
  float dev = 0.f;  // max |x - shift1| over rows with non-zero weight (operand-scale bound for the dense pass)
  double d1[VEC], d2[VEC];
#pragma unroll
  for (int v = 0; v < VEC; ++v) {
    a[v] = (active && shift1) ? shift1[cg * VEC + v] : 0.f;
    b[v] = (active && shift2) ? shift2[cg * VEC + v] : 0.f;
    d1[v] = d2[v] = 0.0;
  }
  if (active) {
    constexpr int UB = 8;  // rows in flight per thread: loads of a batch are issued before any use
    for (int64_t rb = r0 + rl; rb < r1; rb += (int64_t)nrl * 64) {
      float f1[VEC], f2[VEC];
#pragma unroll
      for (int v = 0; v < VEC; ++v) f1[v] = f2[v] = 0.f;
      const int nrows = (int)min((int64_t)64, (r1 - rb + nrl - 1) / nrl);
      for (int it0 = 0; it0 < nrows; it0 += UB) {
        float w[UB], xv[UB][VEC];
#pragma unroll
        for (int u = 0; u < UB; ++u) {
          const bool ok = it0 + u < nrows;
          const int64_t r = ok ? rb + (int64_t)(it0 + u) * nrl : rb;
          w[u] = ok ? __ldg(nw + r) : 0.f;
          if (VEC == 4) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(x + r * D) + cg);
            xv[u][0] = t.x; xv[u][1] = t.y; xv[u][2] = t.z; xv[u][3] = t.w;
          } else {
            xv[u][0] = __ldg(x + r * D + cg);
          }
        }
#pragma unroll
        for (int u = 0; u < UB; ++u) {
          if (w[u] != 0.f) {  // zero-weight rows (NaN rollouts) may hold non-finite samples
#pragma unroll
            for (int v = 0; v < VEC; ++v) {
              const float d0 = xv[u][v] - a[v];
              dev = fmaxf(dev, fabsf(d0));
              f1[v] = fmaf(w[u], d0, f1[v]);
              if (SECOND) {
                const float dv = xv[u][v] - b[v];
                f2[v] = fmaf(w[u] * dv, dv, f2[v]);
              }
            }
          }
        }
      }
#pragma unroll
      for (int v = 0; v < VEC; ++v) {
        d1[v] += (double)f1[v];
        if (SECOND) d2[v] += (double)f2[v];
      }
    }
  }
  if (absdev != nullptr) {
    dev = warp_max(dev);
    if ((threadIdx.x & 31) == 0 && dev > 0.f) atomicMax(absdev, __float_as_uint(dev));
  }
  // combine row lanes
  for (int pass = 0; pass < (SECOND ? 2 : 1); ++pass) {
    __syncthreads();
#pragma unroll
    for (int v = 0; v < VEC; ++v) red[threadIdx.x * VEC + v] = active ? (pass ? d2[v] : d1[v]) : 0.0;
    __syncthreads();
    if (rl == 0 && cg < ncg) {
#pragma unroll
      for (int v = 0; v < VEC; ++v) {
        double t = 0.0;
        for (int q = 0; q < nrl; ++q) t += red[(q * cgb + cg_local) * VEC + v];
        atomicAdd((pass ? s2 : s1) + cg * VEC + v, t);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// acc[i,l] += sum_b nw_b sum_j (x_bij - s_ij)(x_blj - s_lj) inv_j     (H <= 128)
// A group of S samples is staged in shared memory as the H x (S*d) matrix
// A[i, s*d+j] = (x - s) sqrt(nw_b inv_j); acc += A A^T with 4x4 register tiles over the upper
// triangle, the contraction index split across `nlanes` thread groups; fp32 inside a group,
// fp64 across groups, fp64 atomics to global at the end.
// ------------------------------------------------------------------------------------------
template <int MAXB>
__global__ void __launch_bounds__(256)
wmoments_matnorm_kernel(const float* __restrict__ x, const float* __restrict__ nw, int64_t K, int H,
                        int d, const float* __restrict__ shift, const float* __restrict__ inv_cov,
                        double* __restrict__ acc, int S, int Cp, int64_t samples_per_block) {
  extern __shared__ float smem[];
  const int Hb = (H + 3) / 4, Hp = Hb * 4;
  const int nb = Hb * (Hb + 1) / 2;
  float* sA = smem;                 // Hp * Cp  (rows >= H are zero)
  float* sShift = sA + Hp * Cp;     // H*d
  float* sInv = sShift + H * d;     // d
  short* sBi = reinterpret_cast<short*>(sInv + d);  // nb
  short* sBl = sBi + nb;                             // nb

  for (int i = threadIdx.x; i < H * d; i += blockDim.x) sShift[i] = shift ? shift[i] : 0.f;
  for (int i = threadIdx.x; i < d; i += blockDim.x) sInv[i] = inv_cov ? inv_cov[i] : 1.f;
  for (int i = threadIdx.x; i < Hp * Cp; i += blockDim.x) sA[i] = 0.f;
  if (threadIdx.x == 0) {
    int u = 0;
    for (int bi = 0; bi < Hb; ++bi)
      for (int bl = bi; bl < Hb; ++bl) { sBi[u] = (short)bi; sBl[u] = (short)bl; ++u; }
  }
  __syncthreads();

  const int nlanes = (MAXB == 1) ? (int)blockDim.x / nb : 1;
  const int ub0 = (MAXB == 1) ? (int)threadIdx.x % nb : (int)threadIdx.x;
  const int sl = (MAXB == 1) ? (int)threadIdx.x / nb : 0;
  const bool worker = (MAXB == 1) ? (sl < nlanes) : true;

  double dacc[MAXB][16];
#pragma unroll
  for (int m = 0; m < MAXB; ++m)
#pragma unroll
    for (int q = 0; q < 16; ++q) dacc[m][q] = 0.0;

  const int64_t b_begin = (int64_t)blockIdx.x * samples_per_block;
  const int64_t b_end = min(K, b_begin + samples_per_block);
  const int HD = H * d;
  for (int64_t g0 = b_begin; g0 < b_end; g0 += S) {
    const int ns = (int)min((int64_t)S, b_end - g0);
    __syncthreads();
    for (int e = threadIdx.x; e < S * HD; e += blockDim.x) {
      const int s = e / HD, r = e % HD, i = r / d, j = r % d;
      float v = 0.f;
      if (s < ns) {
        const float w = __ldg(nw + g0 + s);
        if (w != 0.f) v = (__ldg(x + (g0 + s) * HD + r) - sShift[r]) * sqrtf(w * sInv[j]);
      }
      sA[i * Cp + s * d + j] = v;
    }
    __syncthreads();
    if (worker) {
      const int C = ns * d;
#pragma unroll
      for (int m = 0; m < MAXB; ++m) {
        const int ub = ub0 + m * (int)blockDim.x;
        if (ub < nb) {
          const float* ra = sA + (sBi[ub] * 4) * Cp;
          const float* rb = sA + (sBl[ub] * 4) * Cp;
          float f[16];
#pragma unroll
          for (int q = 0; q < 16; ++q) f[q] = 0.f;
          for (int c = sl; c < C; c += nlanes) {
            float av[4], bv[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { av[q] = ra[q * Cp + c]; bv[q] = rb[q * Cp + c]; }
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
              for (int q = 0; q < 4; ++q) f[p * 4 + q] = fmaf(av[p], bv[q], f[p * 4 + q]);
          }
#pragma unroll
          for (int q = 0; q < 16; ++q) dacc[m][q] += (double)f[q];
        }
      }
    }
  }
  if (worker) {
#pragma unroll
    for (int m = 0; m < MAXB; ++m) {
      const int ub = ub0 + m * (int)blockDim.x;
      if (ub < nb) {
        const int i0 = sBi[ub] * 4, l0 = sBl[ub] * 4;
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int i = i0 + p, l = l0 + q;
            if (i < H && l < H && l >= i) atomicAdd(acc + (int64_t)i * H + l, dacc[m][p * 4 + q]);
          }
      }
    }
  }
}

__global__ void finalize_gaussian_kernel(const double* __restrict__ s1, const float* __restrict__ cov,
                                         int64_t D, const float* __restrict__ shift, int diag_only,
                                         float* __restrict__ mu_out, float* __restrict__ sigma_out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (mu_out) {
    for (int64_t i = tid; i < D; i += stride)
      mu_out[i] = (float)((shift ? (double)shift[i] : 0.0) + s1[i]);
  }
  if (sigma_out) {
    for (int64_t e = tid; e < D * D; e += stride) {
      const int64_t i = e / D, j = e % D;
      float v;
      if (diag_only) {
        v = (i == j) ? cov[e] : 0.f;
      } else {
        v = 0.5f * (cov[e] + cov[j * D + i]);
      }
      if (s1 && !mu_out) {  // one-pass mode: cov was centred on `shift`, s1 = new mean - shift
        if (!diag_only || i == j) v -= (float)(s1[i] * s1[j]);
      }
      sigma_out[e] = v;
    }
  }
}

__global__ void __launch_bounds__(256)
finalize_matnorm_kernel(const double* __restrict__ s1, const double* __restrict__ acc, int H, int d,
                        double jitter, double* __restrict__ mean64, float* __restrict__ mean32,
                        double* __restrict__ cov64, float* __restrict__ cov32,
                        double* __restrict__ chol64, float* __restrict__ chol32,
                        int* __restrict__ status) {
  extern __shared__ double sa[];
  __shared__ int fail;
  for (int e = threadIdx.x; e < H * d; e += blockDim.x) {
    if (mean64) mean64[e] = s1[e];
    if (mean32) mean32[e] = (float)s1[e];
  }
  if (!acc) return;
  for (int e = threadIdx.x; e < H * H; e += blockDim.x) {
    const int i = e / H, l = e % H;
    const int lo = i < l ? i : l, hi = i < l ? l : i;
    double v = acc[lo * H + hi] / (double)d;  // only the upper triangle was accumulated
    if (i == l) v += jitter;
    if (cov64) cov64[e] = v;
    if (cov32) cov32[e] = (float)v;
    sa[e] = v;
  }
  __syncthreads();
  block_cholesky(sa, H, &fail);
  for (int e = threadIdx.x; e < H * H; e += blockDim.x) {
    const int i = e / H, j = e % H;
    const double v = (j <= i) ? sa[e] : 0.0;
    if (chol64) chol64[e] = v;
    if (chol32) chol32[e] = (float)v;
  }
  if (threadIdx.x == 0 && status) *status = fail;
}


// centre of the covariance pass: the new mean in sample coordinates, fp32
__global__ void __launch_bounds__(256)
matnorm_centre_kernel(const double* __restrict__ s1, const float* __restrict__ shift1, int n,
                      float* __restrict__ shift) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    shift[i] = (float)(s1[i] + (shift1 ? (double)shift1[i] : 0.0));
}

// tail of the packed result: factorisation status, the requested sample row (index given by the host
// or taken from the solver statistics still on the device), and those statistics as raw 8-byte words
__global__ void __launch_bounds__(256)
matnorm_pack_tail_kernel(const int* __restrict__ status, const float* __restrict__ x, int64_t K, int n,
                         int64_t map_index, const ppi_weight_stats* __restrict__ stats,
                         double* __restrict__ tail) {
  if (stats) map_index = stats->argmax;
  const bool have_row = map_index >= 0 && map_index < K;
  if (blockIdx.x == 0 && threadIdx.x == 0) tail[0] = (double)*status;
  if (map_index >= 0 || stats)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
      tail[1 + i] = have_row ? (double)x[map_index * (int64_t)n + i] : nan("");
  if (stats && blockIdx.x == 0 && threadIdx.x < (int)(sizeof(ppi_weight_stats) / 8))
    reinterpret_cast<unsigned long long*>(tail + 1 + n)[threadIdx.x] =
        reinterpret_cast<const unsigned long long*>(stats)[threadIdx.x];
}

}  // namespace ppi

using namespace ppi;

// mma_small.cu
int ppi_wmoments_matnorm_mma(const float* x, const float* nw, int64_t K, int H, int d, const float* shift,
                             const float* inv_cov_out, double* acc, cudaStream_t st);

extern "C" int ppi_wmoments_diag(const float* x, const float* nw, int64_t K, int64_t D,
                                 const float* shift1, const float* shift2, double* s1, double* s2,
                                 float* absdev_max, void* stream) {
  PPI_REQUIRE(x && nw && s1 && K >= 0 && D > 0, PPI_ERR_INVALID, "ppi_wmoments_diag: bad arguments");
  if (K == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
  const int64_t ncg = vec ? D / 4 : D;
  const int cgb = (int)(ncg < 256 ? ncg : 256);
  const int nrl = 256 / cgb;
  const int64_t gx = ceil_div(ncg, cgb);
  // enough blocks for ~8 per SM, each covering at least 64 rows per row lane
  int64_t gy = ceil_div((int64_t)kNumSMs * 8, gx);
  const int64_t max_gy = ceil_div(K, (int64_t)nrl * 64);
  if (gy > max_gy) gy = max_gy;
  if (gy < 1) gy = 1;
  if (gy > 65535) gy = 65535;
  const int64_t rpb = ceil_div(K, gy);
  dim3 grid((unsigned)gx, (unsigned)ceil_div(K, rpb));
  ProfileScope prof(PPI_PROF_WMOMENTS_DIAG, st);
#define PPI_LAUNCH_DIAG(V, SEC) \
  wmoments_diag_kernel<V, SEC><<<grid, 256, 0, st>>>(x, nw, K, D, shift1, shift2, s1, s2,     \
                                                     reinterpret_cast<unsigned int*>(absdev_max), cgb, rpb)
  if (vec) { if (s2) PPI_LAUNCH_DIAG(4, true); else PPI_LAUNCH_DIAG(4, false); }
  else     { if (s2) PPI_LAUNCH_DIAG(1, true); else PPI_LAUNCH_DIAG(1, false); }
#undef PPI_LAUNCH_DIAG
  return check_launch("wmoments_diag_kernel");
}

extern "C" int ppi_wmoments_matnorm(const float* x, const float* nw, int64_t K, int H, int d,
                                    const float* shift, const float* inv_cov_out, double* acc,
                                    void* stream) {
  PPI_REQUIRE(x && nw && acc && K >= 0 && H > 0 && d > 0, PPI_ERR_INVALID,
              "ppi_wmoments_matnorm: bad arguments");
  PPI_REQUIRE(H <= 128, PPI_ERR_UNSUPPORTED, "ppi_wmoments_matnorm supports H <= 128 (got %d)", H);
  if (K == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (H <= 32) {  // warp-level tensor-core kernel (mma_small.cu)
    ProfileScope prof(PPI_PROF_WMOMENTS_MATNORM, st);
    return ppi_wmoments_matnorm_mma(x, nw, K, H, d, shift, inv_cov_out, acc, st);
  }
  const int Hb = (H + 3) / 4, Hp = Hb * 4, nb = Hb * (Hb + 1) / 2;
  // samples per staged group: aim for a contraction length of ~256 within the smem budget
  int S = 256 / d;
  if (S < 1) S = 1;
  if (S > 32) S = 32;
  auto smem_for = [&](int s) {
    const int cp = (s * d) | 1;
    return ((size_t)Hp * cp + (size_t)H * d + d) * sizeof(float) + 2 * (size_t)nb * sizeof(short) + 16;
  };
  while (S > 1 && smem_for(S) > 96 * 1024) S >>= 1;
  PPI_REQUIRE(smem_for(S) <= 200 * 1024, PPI_ERR_UNSUPPORTED,
              "ppi_wmoments_matnorm: H=%d d=%d does not fit shared memory", H, d);
  const int Cp = (S * d) | 1;
  const size_t smem = smem_for(S);
  int64_t grid = ceil_div(K, (int64_t)S * 4);
  if (grid > (int64_t)kNumSMs * 4) grid = (int64_t)kNumSMs * 4;
  if (grid < 1) grid = 1;
  int64_t spb = ceil_div(ceil_div(K, grid), S) * S;
  grid = ceil_div(K, spb);
  ProfileScope prof(PPI_PROF_WMOMENTS_MATNORM, st);
  if (nb <= 256) {
    if (smem > 48 * 1024)
      PPI_CUDA_TRY(cudaFuncSetAttribute(wmoments_matnorm_kernel<1>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    wmoments_matnorm_kernel<1><<<(unsigned)grid, 256, smem, st>>>(x, nw, K, H, d, shift, inv_cov_out,
                                                                  acc, S, Cp, spb);
  } else {
    if (smem > 48 * 1024)
      PPI_CUDA_TRY(cudaFuncSetAttribute(wmoments_matnorm_kernel<3>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    wmoments_matnorm_kernel<3><<<(unsigned)grid, 256, smem, st>>>(x, nw, K, H, d, shift, inv_cov_out,
                                                                  acc, S, Cp, spb);
  }
  return check_launch("wmoments_matnorm_kernel");
}

extern "C" int ppi_finalize_gaussian(const double* s1, const float* cov_in, int64_t D,
                                     const float* shift, int diagonal_only, float* mu_out,
                                     float* sigma_out, void* stream) {
  PPI_REQUIRE(D > 0 && (mu_out || sigma_out), PPI_ERR_INVALID, "ppi_finalize_gaussian: bad arguments");
  PPI_REQUIRE(!mu_out || s1, PPI_ERR_INVALID, "mu_out needs s1");
  PPI_REQUIRE(!sigma_out || cov_in, PPI_ERR_INVALID, "sigma_out needs cov_in");
  PPI_REQUIRE(!sigma_out || cov_in != sigma_out || diagonal_only, PPI_ERR_INVALID,
              "in-place symmetrisation is not supported");
  const int64_t n = sigma_out ? D * D : D;
  const int grid = (int)(ceil_div(n, 256) < kNumSMs * 8 ? ceil_div(n, 256) : kNumSMs * 8);
  finalize_gaussian_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(s1, cov_in, D, shift, diagonal_only,
                                                                   mu_out, sigma_out);
  return check_launch("finalize_gaussian_kernel");
}

extern "C" int ppi_finalize_matnorm(const double* s1, const double* acc, int H, int d, double jitter,
                                    double* mean64, float* mean32, double* cov64, float* cov32,
                                    double* chol64, float* chol32, int* status, void* stream) {
  PPI_REQUIRE(s1 && H > 0 && d > 0, PPI_ERR_INVALID, "ppi_finalize_matnorm: bad arguments");
  size_t smem = 0;
  if (acc) {
    smem = (size_t)H * H * sizeof(double);
    PPI_REQUIRE(smem <= 200 * 1024, PPI_ERR_UNSUPPORTED, "ppi_finalize_matnorm supports H <= 160");
    if (smem > 48 * 1024)
      PPI_CUDA_TRY(cudaFuncSetAttribute(finalize_matnorm_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  finalize_matnorm_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(s1, acc, H, d, jitter, mean64, mean32,
                                                                  cov64, cov32, chol64, chol32, status);
  return check_launch("finalize_matnorm_kernel");
}

// ------------------------------------------------------------------------------------------
// a18-a20 in one call (single rank): weighted mean, covariance about the new mean, symmetrise,
// jitter, Cholesky and the most likely sample, into ONE packed fp64 result the host reads once.
// ------------------------------------------------------------------------------------------
extern "C" size_t ppi_matnorm_update_workspace_bytes(int H, int d) {
  if (H <= 0 || d <= 0) return 0;
  return ((size_t)H * d + (size_t)H * H) * sizeof(double) + ((size_t)H * d + 4) * sizeof(float) + 64;
}

extern "C" int ppi_matnorm_update(const float* x, const float* nw, int64_t K, int H, int d,
                                  const float* shift1, const float* inv_cov_out, double jitter,
                                  int64_t map_index, const ppi_weight_stats* stats, double* packed,
                                  float* mean32, float* chol32, void* workspace, size_t workspace_bytes,
                                  void* stream) {
  static_assert(sizeof(ppi_weight_stats) % 8 == 0, "statistics are copied as 8-byte words");
  PPI_REQUIRE(x && nw && packed && workspace && K > 0 && H > 0 && d > 0, PPI_ERR_INVALID,
              "ppi_matnorm_update: bad arguments");
  PPI_REQUIRE(map_index >= -1 && map_index < K, PPI_ERR_INVALID, "ppi_matnorm_update: map_index out of range");
  PPI_REQUIRE(workspace_bytes >= ppi_matnorm_update_workspace_bytes(H, d), PPI_ERR_WORKSPACE,
              "ppi_matnorm_update: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int HD = H * d;
  double* s1 = (double*)workspace;
  double* acc = s1 + HD;
  float* shift = (float*)(acc + (size_t)H * H);
  int* status = (int*)(shift + HD);
  PPI_CUDA_TRY(cudaMemsetAsync(workspace, 0, ((size_t)HD + (size_t)H * H) * sizeof(double), st));
  PPI_CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(int), st));
  if (int rc = ppi_wmoments_diag(x, nw, K, HD, shift1, nullptr, s1, nullptr, nullptr, stream)) return rc;
  matnorm_centre_kernel<<<(HD + 255) / 256, 256, 0, st>>>(s1, shift1, HD, shift);
  if (int rc = check_launch("matnorm_centre_kernel")) return rc;
  if (int rc = ppi_wmoments_matnorm(x, nw, K, H, d, shift, inv_cov_out, acc, stream)) return rc;
  double* cov = packed + HD;
  double* chol = cov + (size_t)H * H;
  if (int rc = ppi_finalize_matnorm(s1, acc, H, d, jitter, packed, mean32, cov, nullptr, chol, chol32, status,
                                    stream))
    return rc;
  matnorm_pack_tail_kernel<<<(HD + 255) / 256, 256, 0, st>>>(status, x, K, HD, map_index, stats,
                                                             chol + (size_t)H * H);
  return check_launch("matnorm_pack_tail_kernel");
}
