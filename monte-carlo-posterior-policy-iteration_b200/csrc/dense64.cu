// Dense D x D fp64 tail of the Gaussian moment match, hand-written for sm_100a:
//
//   np.linalg.cholesky of the fitted covariance + PD check      ppi/policies.py:259-268
//   log det / trace terms of KL and entropy                      ppi/policies.py:146-158,213-216
//   the factor the next draw samples with                        ppi/samplers.py:38-39
//
// In whitened coordinates (tc_gemm.cu: ppi_wmoments_white) the update needs
//   C = S - m m^T,  G = chol(C),  L' = L G,  mu' = mu + L m,  log det C,  tr C,  |m|^2
// and nothing else: every D^3 term is a product of (block-)triangular fp64 matrices.
//
// Layout.  All matrices row-major.  The factorisation works on 64 x 64 tiles of a padded copy (Dp =
// round_up(D, 64), identity on the padded diagonal).  Right-looking blocked Cholesky, three kernels per block
// column j (the chain  factor -> panel -> trailing update  is the critical path of the whole tail; the
// trailing update is 64-deep rank updates spread over every SM):
//   chol_diag_kernel   one CTA:  L_jj = chol(A_jj) in shared memory, W = L_jj^-1, log det, status
//   chol_panel_kernel  CTA per row block i > j:  X_i = A_ij W^T        (fp64 tensor cores, mma.sync m8n8k4)
//   chol_trail_kernel  CTA per lower tile (i, k), j < k <= i:  A_ik -= X_i X_k^T
// The factor is written twice, G and G^T, so that every product below is an "NT" contraction (both operands
// read along their rows):
//   dgemm_nt_kernel    C_tile(ti, tj) = sum_kt A(ti, kt) B(tj, kt)^T over the k-tiles where both triangular
//                      operands are non-zero (L G: kt in [tj, ti];  L L^T: kt in [0, min(ti, tj)]).
// fp64 MMA fragments (m8n8k4): A (8 x 4) a[lane/4][lane%4], B (4 x 8) b[lane%4][lane/4],
// C (8 x 8) c[lane/4][2 (lane%4) + {0,1}].
#include "common.cuh"

namespace ppi {

constexpr int NB = 64;             // tile edge
constexpr int PITCH = 68;          // shared-memory row pitch in doubles (68 mod 16 = 4: conflict-free fragments)
constexpr int kDenseThreads = 256; // 8 warps: warp (wr, wc) owns rows 32 wr .. +31, columns 16 wc .. +15

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

struct TileAcc {
  double c[4][2][2];   // [row group of 8][column group of 8][2 consecutive columns]
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) c[i][j][0] = c[i][j][1] = 0.0;
  }
};

// acc += As (64 x klen, rows = output rows) * Bs (64 x klen, rows = output columns)^T, both in shared memory
__device__ __forceinline__ void tile_mma(const double* __restrict__ As, const double* __restrict__ Bs,
                                         TileAcc& acc, int klen = NB) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wr = warp >> 2, wc = warp & 3;
  const double* ap = As + (32 * wr + (lane >> 2)) * PITCH + (lane & 3);
  const double* bp = Bs + (16 * wc + (lane >> 2)) * PITCH + (lane & 3);
#pragma unroll 4
  for (int k0 = 0; k0 < klen; k0 += 4) {
    double a[4], b[2];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = ap[i * 8 * PITCH + k0];
#pragma unroll
    for (int j = 0; j < 2; ++j) b[j] = bp[j * 8 * PITCH + k0];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) dmma884(acc.c[i][j], a[i], b[j]);
  }
}

// coordinates of accumulator element (i, j, e) inside the 64 x 64 tile
__device__ __forceinline__ int acc_row(int i) {
  return 32 * ((threadIdx.x >> 5) >> 2) + 8 * i + ((threadIdx.x & 31) >> 2);
}
__device__ __forceinline__ int acc_col(int j) {
  return 16 * ((threadIdx.x >> 5) & 3) + 8 * j + 2 * (threadIdx.x & 3);
}

// 64 x 64 tile of a row-major matrix (leading dimension ld) -> shared memory [64][PITCH]; rows >= rows_valid or
// columns >= cols_valid read as zero (GUARD) -- the padded work matrices need no guard.
template <bool GUARD>
__device__ __forceinline__ void load_tile(double* __restrict__ s, const double* __restrict__ g, int64_t ld,
                                          int rows_valid = NB, int cols_valid = NB) {
  for (int e = threadIdx.x; e < NB * (NB / 2); e += kDenseThreads) {
    const int r = e >> 5, c = (e & 31) * 2;
    double2 v = make_double2(0.0, 0.0);
    if (!GUARD) {
      v = *reinterpret_cast<const double2*>(g + (int64_t)r * ld + c);
    } else if (r < rows_valid) {
      if (c + 1 < cols_valid && ((ld & 1) == 0)) v = *reinterpret_cast<const double2*>(g + (int64_t)r * ld + c);
      else {
        if (c < cols_valid) v.x = g[(int64_t)r * ld + c];
        if (c + 1 < cols_valid) v.y = g[(int64_t)r * ld + c + 1];
      }
    }
    *reinterpret_cast<double2*>(s + r * PITCH + c) = v;
  }
}

// ---------------------------------------------------------------------------------------------
// Cholesky of a padded Dp x Dp matrix held (lower tiles) in `a`; G (lower, zero above the diagonal inside
// the diagonal tiles) and G^T (upper) are written tile by tile.
// ---------------------------------------------------------------------------------------------
struct CholScalars {      // device-side results of one factorisation (fp64, read by the host once)
  double status;          // 0 = positive definite, else 1 + index of the first failing pivot
  double logdet;          // 2 sum log G_ii over the first D pivots
  double trace;           // tr C (white finish only)
  double msq;             // |m|^2 (white finish only)
};

// one CTA: factor the diagonal tile j, invert the factor, accumulate log det.
__global__ void __launch_bounds__(kDenseThreads)
chol_diag_kernel(double* __restrict__ a, int64_t Dp, int j, int64_t D, double* __restrict__ g,
                 double* __restrict__ gt, double* __restrict__ winv, CholScalars* __restrict__ sc) {
  extern __shared__ double smem[];
  double* t = smem;                 // the tile, dense pitch (block_cholesky's layout)
  double* w = smem + NB * NB;       // its inverse, pitch WP (odd: the 4 lanes of a column read 4 banks)
  constexpr int WP = NB + 1;
  __shared__ int fail;
  __shared__ double red[32];
  const int64_t base = (int64_t)j * NB;
  if (sc->status != 0.0) return;    // an earlier block column already failed: nothing below is meaningful
  for (int e = threadIdx.x; e < NB * NB; e += kDenseThreads) {
    const int r = e >> 6, c = e & 63;
    t[e] = a[(base + r) * Dp + base + c];
  }
  block_cholesky(t, NB, &fail);
  if (fail != 0) {
    if (threadIdx.x == 0) sc->status = (double)(base + fail);
    return;
  }
  // W = L^-1, column c by 4 lanes: w_ic = -(sum_{k=c}^{i-1} l_ik w_kc) / l_ii, the dot product split in 4
  {
    const int c = threadIdx.x >> 2, q = threadIdx.x & 3;
    for (int i = 0; i < NB; ++i) {
      double s = 0.0;
      if (i > c)
        for (int k = c + q; k < i; k += 4) s = fma(t[i * NB + k], w[k * WP + c], s);
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      if (q == 0) w[i * WP + c] = i < c ? 0.0 : (i == c ? 1.0 / t[i * NB + i] : -s / t[i * NB + i]);
      __syncwarp();
    }
  }
  __syncthreads();
  double ld = 0.0;
  for (int e = threadIdx.x; e < NB * NB; e += kDenseThreads) {
    const int r = e >> 6, c = e & 63;
    const double v = c <= r ? t[e] : 0.0;
    g[(base + r) * Dp + base + c] = v;
    gt[(base + c) * Dp + base + r] = v;
    winv[e] = w[r * WP + c];
    if (r == c && base + r < D) ld += log(v);
  }
  ld = block_sum(ld, red);
  if (threadIdx.x == 0) sc->logdet += 2.0 * ld;
}

// CTA per row block i = j + 1 + blockIdx.x:  X = A_ij W^T  ->  G(i, j), G^T(j, i)
__global__ void __launch_bounds__(kDenseThreads)
chol_panel_kernel(const double* __restrict__ a, int64_t Dp, int j, const double* __restrict__ winv,
                  double* __restrict__ g, double* __restrict__ gt, const CholScalars* __restrict__ sc) {
  extern __shared__ double smem[];
  double* As = smem;
  double* Bs = smem + NB * PITCH;
  if (sc->status != 0.0) return;
  const int i = j + 1 + blockIdx.x;
  load_tile<false>(As, a + ((int64_t)i * NB) * Dp + (int64_t)j * NB, Dp);
  load_tile<false>(Bs, winv, NB);
  __syncthreads();
  TileAcc acc;
  acc.zero();
  tile_mma(As, Bs, acc);
  __syncthreads();
  // X -> G(i, j) straight from the fragments, and through shared memory (transposed) -> G^T(j, i)
  double* gtile = g + ((int64_t)i * NB) * Dp + (int64_t)j * NB;
#pragma unroll
  for (int ii = 0; ii < 4; ++ii)
#pragma unroll
    for (int jj = 0; jj < 2; ++jj) {
      const int r = acc_row(ii), c = acc_col(jj);
      *reinterpret_cast<double2*>(gtile + (int64_t)r * Dp + c) = make_double2(acc.c[ii][jj][0], acc.c[ii][jj][1]);
      As[c * PITCH + r] = acc.c[ii][jj][0];          // transposed staging: As[col][row]
      As[(c + 1) * PITCH + r] = acc.c[ii][jj][1];
    }
  __syncthreads();
  double* gttile = gt + ((int64_t)j * NB) * Dp + (int64_t)i * NB;
  for (int e = threadIdx.x; e < NB * (NB / 2); e += kDenseThreads) {
    const int r = e >> 5, c = (e & 31) * 2;
    *reinterpret_cast<double2*>(gttile + (int64_t)r * Dp + c) = *reinterpret_cast<const double2*>(As + r * PITCH + c);
  }
}

// CTA per lower tile (i, k) of the trailing matrix, j < k <= i:  A_ik -= X_i X_k^T
__global__ void __launch_bounds__(kDenseThreads)
chol_trail_kernel(double* __restrict__ a, int64_t Dp, int j, int T, const double* __restrict__ g,
                  const CholScalars* __restrict__ sc) {
  extern __shared__ double smem[];
  double* As = smem;
  double* Bs = smem + NB * PITCH;
  if (sc->status != 0.0) return;
  // blockIdx.x -> (i, k) with j < k <= i < T: row-major enumeration of the lower triangle of size n = T - j - 1
  int rem = blockIdx.x, ri = 0;
  while (rem > ri) { rem -= ri + 1; ++ri; }
  const int i = j + 1 + ri, k = j + 1 + rem;
  load_tile<false>(As, g + ((int64_t)i * NB) * Dp + (int64_t)j * NB, Dp);
  load_tile<false>(Bs, g + ((int64_t)k * NB) * Dp + (int64_t)j * NB, Dp);
  __syncthreads();
  TileAcc acc;
  acc.zero();
  tile_mma(As, Bs, acc);
  double* ct = a + ((int64_t)i * NB) * Dp + (int64_t)k * NB;
#pragma unroll
  for (int ii = 0; ii < 4; ++ii)
#pragma unroll
    for (int jj = 0; jj < 2; ++jj) {
      double2* p = reinterpret_cast<double2*>(ct + (int64_t)acc_row(ii) * Dp + acc_col(jj));
      double2 v = *p;
      v.x -= acc.c[ii][jj][0];
      v.y -= acc.c[ii][jj][1];
      *p = v;
    }
}

// ---------------------------------------------------------------------------------------------
// C(ti, tj) = sum_kt A(ti, kt) B(tj, kt)^T over k-tiles [klo, khi]; tiles enumerated heaviest first.
//   MODE_LG   A = L (lower), B = G^T (upper, padded Dp): lower tiles, kt in [tj, ti]
//   MODE_LLT  A = B = L (lower):                         lower tiles, kt in [0, tj]; mirrored into the upper part
// A is read with leading dimension lda and guards (D x D, any D); B = padded work matrix or A itself.
// ---------------------------------------------------------------------------------------------
enum { MODE_LG = 0, MODE_LLT = 1 };

template <int MODE>
__global__ void __launch_bounds__(kDenseThreads)
dgemm_nt_kernel(const double* __restrict__ A, int64_t lda, const double* __restrict__ B, int64_t ldb, int64_t D,
                int T, double* __restrict__ out64, float* __restrict__ out32, int64_t ldo,
                const CholScalars* __restrict__ sc) {
  extern __shared__ double smem[];
  double* As = smem;
  double* Bs = smem + NB * PITCH;
  if (sc && sc->status != 0.0) return;
  // heaviest tiles first.  LG: weight = ti - tj (offset d): d = T-1 (1 tile), T-2 (2 tiles), ...
  // LLT: weight = tj: column tj = T-1 (1 tile), T-2 (2 tiles), ...
  int rem = blockIdx.x, cnt = 1;
  while (rem >= cnt) { rem -= cnt; ++cnt; }
  int ti, tj;
  if (MODE == MODE_LG) { const int d = T - cnt; tj = rem; ti = tj + d; }
  else { tj = T - cnt; ti = tj + rem; }
  const int klo = MODE == MODE_LG ? tj : 0, khi = MODE == MODE_LG ? ti : tj;
  const int rows_a = (int)min((int64_t)NB, D - (int64_t)ti * NB);
  const int rows_b = (int)min((int64_t)NB, D - (int64_t)tj * NB);
  TileAcc acc;
  acc.zero();
  for (int kt = klo; kt <= khi; ++kt) {
    const int cols = (int)min((int64_t)NB, D - (int64_t)kt * NB);
    load_tile<true>(As, A + ((int64_t)ti * NB) * lda + (int64_t)kt * NB, lda, rows_a, cols);
    if (MODE == MODE_LG) load_tile<false>(Bs, B + ((int64_t)tj * NB) * ldb + (int64_t)kt * NB, ldb);
    else load_tile<true>(Bs, B + ((int64_t)tj * NB) * ldb + (int64_t)kt * NB, ldb, rows_b, cols);
    __syncthreads();
    tile_mma(As, Bs, acc);
    __syncthreads();
  }
#pragma unroll
  for (int ii = 0; ii < 4; ++ii)
#pragma unroll
    for (int jj = 0; jj < 2; ++jj) {
      const int64_t r = (int64_t)ti * NB + acc_row(ii), c = (int64_t)tj * NB + acc_col(jj);
      if (r >= D) continue;
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        if (c + e >= D) continue;
        const double v = acc.c[ii][jj][e];
        if (out64) out64[r * ldo + c + e] = v;
        if (out32) out32[r * ldo + c + e] = (float)v;
        if (MODE == MODE_LLT && ti != tj) {
          if (out64) out64[(c + e) * ldo + r] = v;
          if (out32) out32[(c + e) * ldo + r] = (float)v;
        }
      }
    }
}

// ---------------------------------------------------------------------------------------------
// small element-wise kernels around the factorisation
// ---------------------------------------------------------------------------------------------
// work = padded copy of the lower triangle of src (D x D fp64, ld = D); identity on the padded diagonal
__global__ void __launch_bounds__(256)
pad_lower_kernel(const double* __restrict__ src, int64_t D, int64_t Dp, double* __restrict__ work) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < Dp * Dp; e += stride) {
    const int64_t i = e / Dp, j = e % Dp;
    if (j > i) continue;
    work[e] = (i < D && j < D) ? src[i * D + j] : (i == j ? 1.0 : 0.0);
  }
}

// work = S - m m^T (lower triangle, padded), trace and |m|^2 into the scalars
__global__ void __launch_bounds__(256)
white_c_kernel(const float* __restrict__ S, const double* __restrict__ m, const double* __restrict__ diag,
               int64_t D, int64_t Dp, double* __restrict__ work, CholScalars* __restrict__ sc) {
  __shared__ double red[32];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  double tr = 0.0, msq = 0.0;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < Dp * Dp; e += stride) {
    const int64_t i = e / Dp, j = e % Dp;
    if (j > i) continue;
    double v = i == j ? 1.0 : 0.0;
    if (i < D && j < D) {
      const double mi = m[i], mj = m[j];
      v = ((i == j && diag) ? diag[i] : (double)S[i * D + j]) - mi * mj;
      if (i == j) { tr += v; msq += mi * mi; }
    }
    work[e] = v;
  }
  tr = block_sum(tr, red);
  msq = block_sum(msq, red);
  if (threadIdx.x == 0 && (tr != 0.0 || msq != 0.0)) {
    atomicAdd(&sc->trace, tr);
    atomicAdd(&sc->msq, msq);
  }
}

// L (D x D, ld = D, zero upper triangle) and optionally its transpose, from the padded factor
__global__ void __launch_bounds__(256)
unpad_lower_kernel(const double* __restrict__ g, int64_t D, int64_t Dp, double* __restrict__ L64,
                   float* __restrict__ L32) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < D * D; e += stride) {
    const int64_t i = e / D, j = e % D;
    const double v = j <= i ? g[i * Dp + j] : 0.0;
    if (L64) L64[e] = v;
    if (L32) L32[e] = (float)v;
  }
}

// y = x0 + L v (L lower, D x D fp64): one warp per row
__global__ void __launch_bounds__(256)
tri_matvec_kernel(const double* __restrict__ L, int64_t D, const double* __restrict__ v,
                  const float* __restrict__ x0, float* __restrict__ y32, double* __restrict__ y64) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= D) return;
  double s = 0.0;
  for (int64_t k = lane; k <= row; k += 32) s = fma(L[row * D + k], v[k], s);
  s = warp_sum(s);
  if (lane == 0) {
    const double r = (x0 ? (double)x0[row] : 0.0) + s;
    if (y32) y32[row] = (float)r;
    if (y64) y64[row] = r;
  }
}

constexpr size_t kTileSmem = 2 * NB * PITCH * sizeof(double);   // 69 632 B: opt-in above 48 KB
constexpr size_t kDiagSmem = (size_t)(NB * NB + NB * (NB + 1)) * sizeof(double);

static int64_t pad64(int64_t D) { return (D + NB - 1) / NB * NB; }

static int set_dense_attrs() {
  static bool done = false;
  if (done) return 0;
  PPI_CUDA_TRY(cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDiagSmem));
  PPI_CUDA_TRY(cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTileSmem));
  PPI_CUDA_TRY(cudaFuncSetAttribute(chol_trail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTileSmem));
  PPI_CUDA_TRY(cudaFuncSetAttribute(dgemm_nt_kernel<MODE_LG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTileSmem));
  PPI_CUDA_TRY(cudaFuncSetAttribute(dgemm_nt_kernel<MODE_LLT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTileSmem));
  done = true;
  return 0;
}

// workspace: [scalars 1 KB][work Dp^2][G Dp^2][G^T Dp^2][W 64^2]
struct DenseWs {
  CholScalars* sc;
  double *work, *g, *gt, *winv;
};
static size_t dense_ws_bytes(int64_t D) {
  const int64_t Dp = pad64(D);
  return 1024 + (size_t)(3 * Dp * Dp + NB * NB) * sizeof(double) + 256;
}
static DenseWs dense_ws(void* workspace, int64_t D) {
  const int64_t Dp = pad64(D);
  char* p = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  DenseWs w;
  w.sc = (CholScalars*)p;
  w.work = (double*)(p + 1024);
  w.g = w.work + Dp * Dp;
  w.gt = w.g + Dp * Dp;
  w.winv = w.gt + Dp * Dp;
  return w;
}

// blocked factorisation of the padded matrix in ws.work -> ws.g / ws.gt, scalars in ws.sc
static int chol_blocked(const DenseWs& ws, int64_t D, cudaStream_t st) {
  const int64_t Dp = pad64(D);
  const int T = (int)(Dp / NB);
  for (int j = 0; j < T; ++j) {
    chol_diag_kernel<<<1, kDenseThreads, kDiagSmem, st>>>(ws.work, Dp, j, D, ws.g, ws.gt, ws.winv, ws.sc);
    const int below = T - j - 1;
    if (below > 0) {
      chol_panel_kernel<<<below, kDenseThreads, kTileSmem, st>>>(ws.work, Dp, j, ws.winv, ws.g, ws.gt, ws.sc);
      chol_trail_kernel<<<below * (below + 1) / 2, kDenseThreads, kTileSmem, st>>>(ws.work, Dp, j, T, ws.g, ws.sc);
    }
  }
  return check_launch("blocked Cholesky kernels", 3 * T - 2);
}

}  // namespace ppi

using namespace ppi;

extern "C" size_t ppi_dense_workspace_bytes(int64_t D) { return D > 0 ? dense_ws_bytes(D) : 0; }

extern "C" int ppi_cholesky_blocked(const double* a, int64_t D, double* chol64, float* chol32, double* scalars,
                                    void* workspace, size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(a && D > 0 && scalars && workspace && (chol64 || chol32), PPI_ERR_INVALID,
              "ppi_cholesky_blocked: bad arguments");
  PPI_REQUIRE(workspace_bytes >= dense_ws_bytes(D), PPI_ERR_WORKSPACE, "ppi_cholesky_blocked: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = set_dense_attrs()) return rc;
  const DenseWs ws = dense_ws(workspace, D);
  const int64_t Dp = pad64(D);
  ProfileScope prof(PPI_PROF_DENSE_TAIL, st);
  PPI_CUDA_TRY(cudaMemsetAsync(ws.sc, 0, sizeof(CholScalars), st));
  pad_lower_kernel<<<(unsigned)min((int64_t)kNumSMs * 8, ceil_div(Dp * Dp, 256)), 256, 0, st>>>(a, D, Dp, ws.work);
  if (int rc = check_launch("pad_lower_kernel")) return rc;
  if (int rc = chol_blocked(ws, D, st)) return rc;
  unpad_lower_kernel<<<(unsigned)min((int64_t)kNumSMs * 8, ceil_div(D * D, 256)), 256, 0, st>>>(ws.g, D, Dp, chol64,
                                                                                                  chol32);
  PPI_CUDA_TRY(cudaMemcpyAsync(scalars, ws.sc, sizeof(CholScalars), cudaMemcpyDeviceToDevice, st));
  return check_launch("unpad_lower_kernel");
}

extern "C" int ppi_white_finish(const float* S, const double* m, const double* diag, const double* l_old,
                                const float* mu_old, int64_t D, double* l_new64, float* l_new32, float* mu_new32,
                                double* scalars, void* workspace, size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(S && m && l_old && mu_old && D > 0 && l_new64 && l_new32 && mu_new32 && scalars && workspace,
              PPI_ERR_INVALID, "ppi_white_finish: bad arguments");
  PPI_REQUIRE(workspace_bytes >= dense_ws_bytes(D), PPI_ERR_WORKSPACE, "ppi_white_finish: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = set_dense_attrs()) return rc;
  const DenseWs ws = dense_ws(workspace, D);
  const int64_t Dp = pad64(D);
  const int T = (int)(Dp / NB);
  ProfileScope prof(PPI_PROF_DENSE_TAIL, st);
  PPI_CUDA_TRY(cudaMemsetAsync(ws.sc, 0, sizeof(CholScalars), st));
  white_c_kernel<<<(unsigned)min((int64_t)kNumSMs * 8, ceil_div(Dp * Dp, 256)), 256, 0, st>>>(S, m, diag, D, Dp, ws.work, ws.sc);
  tri_matvec_kernel<<<(unsigned)ceil_div(D * 32, 256), 256, 0, st>>>(l_old, D, m, mu_old, mu_new32, nullptr);
  if (int rc = check_launch("white_c_kernel / tri_matvec_kernel", 2)) return rc;
  if (int rc = chol_blocked(ws, D, st)) return rc;
  dgemm_nt_kernel<MODE_LG><<<T * (T + 1) / 2, kDenseThreads, kTileSmem, st>>>(l_old, D, ws.gt, Dp, D, T, l_new64,
                                                                              l_new32, D, ws.sc);
  PPI_CUDA_TRY(cudaMemcpyAsync(scalars, ws.sc, sizeof(CholScalars), cudaMemcpyDeviceToDevice, st));
  return check_launch("dgemm_nt_kernel<LG>");
}

extern "C" int ppi_tri_gram(const double* l, int64_t D, double* out64, float* out32, void* stream) {
  PPI_REQUIRE(l && D > 0 && (out64 || out32), PPI_ERR_INVALID, "ppi_tri_gram: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = set_dense_attrs()) return rc;
  const int T = (int)(pad64(D) / NB);
  ProfileScope prof(PPI_PROF_DENSE_TAIL, st);
  dgemm_nt_kernel<MODE_LLT><<<T * (T + 1) / 2, kDenseThreads, kTileSmem, st>>>(l, D, l, D, D, T, out64, out32, D,
                                                                               nullptr);
  return check_launch("dgemm_nt_kernel<LLT>");
}

extern "C" int ppi_tri_matvec(const double* l, int64_t D, const double* v, const float* x0, float* y32,
                              void* stream) {
  PPI_REQUIRE(l && v && y32 && D > 0, PPI_ERR_INVALID, "ppi_tri_matvec: bad arguments");
  tri_matvec_kernel<<<(unsigned)ceil_div(D * 32, 256), 256, 0, (cudaStream_t)stream>>>(l, D, v, x0, y32, nullptr);
  return check_launch("tri_matvec_kernel");
}
