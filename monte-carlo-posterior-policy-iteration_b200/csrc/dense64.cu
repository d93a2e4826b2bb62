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
// round_up(D, 64), identity on the padded diagonal).  Right-looking blocked Cholesky with the look-ahead built
// into the launch structure -- ONE launch per block column j (chol_step_kernel), whose CTAs never wait on one
// another:
//   A-role (CTA per row block i >= j): apply panel j-1's pending update to the own tiles of block column j
//     (fp64 tensor cores, mma.sync m8n8k4), factor the diagonal tile in registers (4 x 4 blocks per thread; every
//     CTA redundantly), solve the panel tile  X_i = A_ij L_jj^-T  by blocked forward substitution;
//   B-role (CTA per lower tile (i, k), k > j): the rest of panel j-1's trailing update  A_ik -= X_i X_k^T.
// The critical path is a chain of D / 64 launches of ~one tile factorisation + one tile solve each.
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

// The same tile through cp.async (LDGSTS, 16 bytes per request, no registers in between): every thread has its
// eight requests per tile in flight at once, and several tiles can be queued before one wait -- a plain
// load / store loop exposes one L2 round trip per iteration (measured: 14 k cycles for the four tiles of a step).
__device__ __forceinline__ void load_tile_async(double* __restrict__ s, const double* __restrict__ g, int64_t ld) {
#pragma unroll
  for (int it = 0; it < NB * (NB / 2) / kDenseThreads; ++it) {
    const int e = threadIdx.x + it * kDenseThreads;
    const int r = e >> 5, c = (e & 31) * 2;
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s + r * PITCH + c);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(g + (int64_t)r * ld + c) : "memory");
  }
}
__device__ __forceinline__ void async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

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

// ---- SIMT part of a step: 256 threads = 16 x 16, thread (ty, tx) owns the 4 x 4 block (rows 4 ty.., columns
// 4 tx..) of a 64 x 64 tile in registers.  Blocks of the factor are exchanged through shared memory in the layout
//   Lt[kb][(a * 4 + c) * 16 + t]  =  element (a, c) of block (t, kb)          (t = block row, kb = block column)
// so that the 16 threads of a block row / column read consecutive doubles (no bank conflicts) and the
// threads that share a block read one address (broadcast).
constexpr int kLtBlock = 256;     // doubles per block column of Lt: 16 entries x 16 block rows

// 4 x 4 Cholesky of diagonal block kb (held in r by its owner thread) and the inverse of its factor, unrolled in
// registers; publishes the factor (Lt) and its inverse (dinv).  One reciprocal square root per pivot
// (l = p rsqrt(p), 1 / l = rsqrt(p)): fp64 sqrt and division are long dependent software sequences, and this
// chain is the serial spine of the whole tail.
__device__ __forceinline__ void factor_diag_block(double (&r)[4][4], int kb, double* __restrict__ Lt,
                                                  double* __restrict__ dinv, int* fail_s) {
  int bad = 0;
  auto ok = [](double p) { return p > 0.0 && isfinite(p); };
  const double p0 = r[0][0];
  if (!ok(p0)) bad = 1;
  const double i0 = rsqrt(p0), l00 = p0 * i0;
  const double l10 = r[1][0] * i0, l20 = r[2][0] * i0, l30 = r[3][0] * i0;
  const double p1 = r[1][1] - l10 * l10;
  if (!bad && !ok(p1)) bad = 2;
  const double i1 = rsqrt(p1), l11 = p1 * i1;
  const double l21 = (r[2][1] - l20 * l10) * i1, l31 = (r[3][1] - l30 * l10) * i1;
  const double p2 = r[2][2] - l20 * l20 - l21 * l21;
  if (!bad && !ok(p2)) bad = 3;
  const double i2 = rsqrt(p2), l22 = p2 * i2;
  const double l32 = (r[3][2] - l30 * l20 - l31 * l21) * i2;
  const double p3 = r[3][3] - l30 * l30 - l31 * l31 - l32 * l32;
  if (!bad && !ok(p3)) bad = 4;
  const double i3 = rsqrt(p3), l33 = p3 * i3;
  if (bad) *fail_s = 4 * kb + bad;
  const double w10 = -l10 * i0 * i1, w21 = -l21 * i1 * i2, w32 = -l32 * i2 * i3;
  const double w20 = -(l20 * i0 + l21 * w10) * i2, w31 = -(l31 * i1 + l32 * w21) * i3;
  const double w30 = -(l30 * i0 + l31 * w10 + l32 * w20) * i3;
  double* w = dinv + kb * 16;
  w[0] = i0;  w[1] = 0.0; w[2] = 0.0;  w[3] = 0.0;
  w[4] = w10; w[5] = i1;  w[6] = 0.0;  w[7] = 0.0;
  w[8] = w20; w[9] = w21; w[10] = i2;  w[11] = 0.0;
  w[12] = w30; w[13] = w31; w[14] = w32; w[15] = i3;
  r[0][0] = l00; r[0][1] = 0.0; r[0][2] = 0.0; r[0][3] = 0.0;
  r[1][0] = l10; r[1][1] = l11; r[1][2] = 0.0; r[1][3] = 0.0;
  r[2][0] = l20; r[2][1] = l21; r[2][2] = l22; r[2][3] = 0.0;
  r[3][0] = l30; r[3][1] = l31; r[3][2] = l32; r[3][3] = l33;
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int c = 0; c < 4; ++c) Lt[kb * kLtBlock + (a * 4 + c) * 16 + kb] = r[a][c];
}

// Right-looking Cholesky of the 64 x 64 tile in Ds (pitch PITCH, lower part).  On return thread (ty >= tx) holds
// block (ty, tx) of the factor in r, Lt holds every block column, dinv[kb] the inverse of diagonal block kb.
// Returns false (uniformly) when a pivot fails; *fail_s = 1 + index of the first failing pivot in the tile.
// Two barriers per block column; the owner of the NEXT diagonal block factors it right after its own trailing
// update, while the other threads are still updating theirs (the scalar chain hides behind the rank-4 update).
__device__ __forceinline__ bool tile_cholesky(const double* __restrict__ Ds, double* __restrict__ Lt,
                                              double* __restrict__ dinv, int* fail_s, double (&r)[4][4]) {
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  const bool lower = ty >= tx;
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int c = 0; c < 4; ++c) r[a][c] = lower ? Ds[(4 * ty + a) * PITCH + 4 * tx + c] : 0.0;
  if (ty == 0 && tx == 0) factor_diag_block(r, 0, Lt, dinv, fail_s);
  for (int kb = 0; kb < 16; ++kb) {
    __syncthreads();                        // diagonal block kb factored, its inverse published
    if (*fail_s != 0) return false;
    if (tx == kb && ty > kb) {
      // X = A W^T with W = inverse of the diagonal factor: the final block (ty, kb) of the factor
      const double* w = dinv + kb * 16;
      double x[4][4];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b2 = 0; b2 < 4; ++b2) {
          double acc = 0.0;
#pragma unroll
          for (int c = 0; c <= b2; ++c) acc = fma(r[a][c], w[b2 * 4 + c], acc);
          x[a][b2] = acc;
        }
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          r[a][c] = x[a][c];
          Lt[kb * kLtBlock + (a * 4 + c) * 16 + ty] = x[a][c];
        }
    }
    __syncthreads();                        // block column kb published
    if (tx > kb && lower) {
      const double* pl = Lt + kb * kLtBlock;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        double pa[4], pb[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { pa[a] = pl[(a * 4 + c) * 16 + ty]; pb[a] = pl[(a * 4 + c) * 16 + tx]; }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b2 = 0; b2 < 4; ++b2) r[a][b2] = fma(-pa[a], pb[b2], r[a][b2]);
      }
      if (ty == kb + 1 && tx == kb + 1) factor_diag_block(r, kb + 1, Lt, dinv, fail_s);
    }
  }
  __syncthreads();
  return *fail_s == 0;
}

// ---- panel solve on fp64 tensor cores ------------------------------------------------------------------------
// X = T L^-T row by row is independent across rows, so every warp solves ITS 8 rows of the tile alone (no block
// barrier): four 16-column blocks,  R_b = T_b - sum_{p<b} X_p L_bp^T  (m8n8k4 chain of 4 b k-steps),
// X_b = R_b W_b^T with W_b the inverse of the b-th 16 x 16 diagonal block of L.  ~80 DMMA per warp instead of
// sixteen block-wide SIMT phases (measured 16 k cycles).
constexpr int WP16 = 20;   // pitch of the 16 x 16 inverses and of the per-warp R scratch (20 mod 16 = 4: conflict-free)

// Inverses of the four 16 x 16 diagonal blocks of the factor from its 4 x 4 blocks (Lt) and their inverses (dinv):
// W_aa = dinv,  W_ab = -W_aa sum_{k=b}^{a-1} L_ak W_kb  (a > b), three dependency levels inside warp 0.
__device__ __forceinline__ void build_w16(const double* __restrict__ Lt, const double* __restrict__ dinv,
                                          double* __restrict__ w16) {
  const int tid = threadIdx.x;
  // diagonal 4 x 4 blocks and zeros everywhere else: 4 blocks x 16 x 16 entries
  for (int e = tid; e < 4 * 16 * 16; e += kDenseThreads) {
    const int blk = e >> 8, rr = (e >> 4) & 15, cc = e & 15;
    double v = 0.0;
    if ((rr >> 2) == (cc >> 2)) v = dinv[(4 * blk + (rr >> 2)) * 16 + (rr & 3) * 4 + (cc & 3)];
    w16[blk * 16 * WP16 + rr * WP16 + cc] = v;
  }
  __syncthreads();
  if (tid < 32) {
    const int blk = tid / 6, e = tid % 6;   // (a, b) = (1,0) (2,1) (3,2) | (2,0) (3,1) | (3,0)
    const int a = e < 3 ? e + 1 : (e < 5 ? e - 1 : 3), b = e < 3 ? e : (e < 5 ? e - 3 : 0);
    double* wb = w16 + blk * 16 * WP16;
#pragma unroll
    for (int d = 1; d <= 3; ++d) {
      if (tid < 24 && a - b == d) {
        double sacc[4][4];
#pragma unroll
        for (int x = 0; x < 4; ++x)
#pragma unroll
          for (int y = 0; y < 4; ++y) sacc[x][y] = 0.0;
        for (int k = b; k < a; ++k) {
          // L block (row 4 blk + a, column 4 blk + k) times W block (k, b)
          const double* lb = Lt + (4 * blk + k) * kLtBlock + (4 * blk + a);
#pragma unroll
          for (int x = 0; x < 4; ++x)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const double l = lb[(x * 4 + c) * 16];
#pragma unroll
              for (int y = 0; y < 4; ++y) sacc[x][y] = fma(l, wb[(4 * k + c) * WP16 + 4 * b + y], sacc[x][y]);
            }
        }
        const double* waa = dinv + (4 * blk + a) * 16;
#pragma unroll
        for (int x = 0; x < 4; ++x)
#pragma unroll
          for (int y = 0; y < 4; ++y) {
            double v = 0.0;
#pragma unroll
            for (int c = 0; c <= x; ++c) v = fma(waa[x * 4 + c], sacc[c][y], v);
            wb[(4 * a + x) * WP16 + 4 * b + y] = -v;
          }
      }
      __syncwarp();
    }
  }
  __syncthreads();
}

// In place: Ts (64 x 64, pitch PITCH) <- Ts L^-T, with L row-major in Ls (pitch PITCH, zero above the diagonal),
// the 16 x 16 inverses in w16 and 8 x (8 x WP16) doubles of per-warp scratch in rs.
__device__ __forceinline__ void tile_solve_mma(double* __restrict__ Ts, const double* __restrict__ Ls,
                                               const double* __restrict__ w16, double* __restrict__ rs,
                                               double* __restrict__ gtile, double* __restrict__ gttile, int64_t Dp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int fr = lane >> 2, fk = lane & 3;           // fragment row / contraction index
  double* trow = Ts + (8 * warp) * PITCH;            // this warp's 8 rows
  double* rw = rs + warp * 8 * WP16;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    double acc[2][2];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      acc[nt][0] = trow[fr * PITCH + 16 * b + 8 * nt + 2 * fk];
      acc[nt][1] = trow[fr * PITCH + 16 * b + 8 * nt + 2 * fk + 1];
    }
    // R_b = T_b - X[:, 0 : 16 b] L[16 b .. 16 b + 15, 0 : 16 b]^T
    for (int k0 = 0; k0 < 16 * b; k0 += 4) {
      const double av = -trow[fr * PITCH + k0 + fk];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) dmma884(acc[nt], av, Ls[(16 * b + 8 * nt + fr) * PITCH + k0 + fk]);
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      rw[fr * WP16 + 8 * nt + 2 * fk] = acc[nt][0];
      rw[fr * WP16 + 8 * nt + 2 * fk + 1] = acc[nt][1];
    }
    __syncwarp();
    // X_b = R_b W_b^T
    double xacc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
    const double* wb = w16 + b * 16 * WP16;
#pragma unroll
    for (int k0 = 0; k0 < 16; k0 += 4) {
      const double av = rw[fr * WP16 + k0 + fk];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) dmma884(xacc[nt], av, wb[(8 * nt + fr) * WP16 + k0 + fk]);
    }
    // X_b: back into the tile (operand of the later blocks) and straight from the fragments to G(i, j) (16 bytes
    // per lane, 64 contiguous bytes per row) and G^T(j, i) (8 consecutive rows = 64 contiguous bytes per column)
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      const int col = 16 * b + 8 * nt + 2 * fk, row = 8 * warp + fr;
      trow[fr * PITCH + col] = xacc[nt][0];
      trow[fr * PITCH + col + 1] = xacc[nt][1];
      *reinterpret_cast<double2*>(gtile + (int64_t)row * Dp + col) = make_double2(xacc[nt][0], xacc[nt][1]);
      gttile[(int64_t)col * Dp + row] = xacc[nt][0];
      gttile[(int64_t)(col + 1) * Dp + row] = xacc[nt][1];
    }
    __syncwarp();
  }
}

// subtract the DMMA accumulator from a tile in shared memory (each element has one owner)
__device__ __forceinline__ void tile_sub(double* __restrict__ s, const TileAcc& acc) {
#pragma unroll
  for (int ii = 0; ii < 4; ++ii)
#pragma unroll
    for (int jj = 0; jj < 2; ++jj) {
      double2* p = reinterpret_cast<double2*>(s + acc_row(ii) * PITCH + acc_col(jj));
      double2 v = *p;
      v.x -= acc.c[ii][jj][0];
      v.y -= acc.c[ii][jj][1];
      *p = v;
    }
}

// Phase clocks of the critical path (A-role CTAs i = j and i = j + 1 of every step), read by tools/dense_tail_timing.py
// through ppi_debug_chol_trace: [step][cta 0/1][phase] SM clock at start / loaded / updated / factored / done.
__device__ long long g_chol_trace[64 * 2 * 8];
#define CHOL_TRACE(slot) do { if (trace_me && threadIdx.x == 0) g_chol_trace[((j & 63) * 2 + (int)blockIdx.x) * 8 + (slot)] = clock64(); } while (0)

// One launch per block column j (look-ahead built in: nothing inside a launch waits on another CTA).
//   A-role, CTA per row block i = j .. T-1: applies the PENDING update of panel j-1 to its own tiles of block
//     column j (A_jj -= X_j X_j^T, A_ij -= X_i X_j^T with X = G(., j-1): fp64 tensor cores), factors the
//     diagonal tile (every CTA redundantly: the other SMs would idle), then either writes the diagonal factor
//     (i == j) or solves its panel tile X_i = A_ij L_jj^-T and writes G(i, j), G^T(j, i);
//   B-role, CTA per lower tile (i, k), j+1 <= k <= i: the rest of panel j-1's trailing update,
//     A_ik -= X_i X_k^T -- independent of this step's factorisation, so it runs beside it.
// The critical path of the factorisation is one launch per block column.
__global__ void __launch_bounds__(kDenseThreads, 1)
chol_step_kernel(double* __restrict__ a, int64_t Dp, int j, int T, int64_t D, double* __restrict__ g,
                 double* __restrict__ gt, CholScalars* __restrict__ sc) {
  extern __shared__ double smem[];
  double* As = smem;                       // X_i           (MMA operand)
  double* Bs = As + NB * PITCH;            // X_j / X_k     (MMA operand)
  double* Ds = Bs + NB * PITCH;            // diagonal tile of block column j
  double* Ts = Ds + NB * PITCH;            // this CTA's panel tile
  double* Lt = Ts + NB * PITCH;            // factor blocks, 16 x 256
  double* dinv = Lt + 16 * kLtBlock;       // 16 inverses of the diagonal 4 x 4 blocks
  double* w16 = dinv + 256;                // inverses of the four 16 x 16 diagonal blocks, pitch WP16
  double* rs = w16 + 4 * 16 * WP16;        // per-warp 8 x 16 scratch of tile_solve_mma
  __shared__ int fail;
  __shared__ double red[32];
  if (sc->status != 0.0) return;           // an earlier block column failed
  const int nA = T - j;
  if ((int)blockIdx.x >= nA) {
    // ---------------- B-role ----------------
    int rem = blockIdx.x - nA, ri = 0;
    while (rem > ri) { rem -= ri + 1; ++ri; }
    const int i = j + 1 + ri, k = j + 1 + rem;
    load_tile_async(As, g + ((int64_t)i * NB) * Dp + (int64_t)(j - 1) * NB, Dp);
    load_tile_async(Bs, g + ((int64_t)k * NB) * Dp + (int64_t)(j - 1) * NB, Dp);
    async_commit();
    async_wait<0>();
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
    return;
  }
  // ---------------- A-role ----------------
  const int i = j + blockIdx.x;
  const bool diag_cta = i == j;
  const bool trace_me = blockIdx.x < 2;
  CHOL_TRACE(0);
  if (threadIdx.x == 0) fail = 0;
  // two request groups: what the diagonal tile needs first, the panel tile's operands behind it
  load_tile_async(Ds, a + ((int64_t)j * NB) * Dp + (int64_t)j * NB, Dp);
  if (j > 0) load_tile_async(Bs, g + ((int64_t)j * NB) * Dp + (int64_t)(j - 1) * NB, Dp);
  async_commit();
  if (!diag_cta) {
    load_tile_async(Ts, a + ((int64_t)i * NB) * Dp + (int64_t)j * NB, Dp);
    if (j > 0) load_tile_async(As, g + ((int64_t)i * NB) * Dp + (int64_t)(j - 1) * NB, Dp);
  }
  async_commit();
  if (j > 0) {
    async_wait<1>();
    __syncthreads();
    CHOL_TRACE(1);
    TileAcc acc;
    acc.zero();
    tile_mma(Bs, Bs, acc);
    tile_sub(Ds, acc);
    if (!diag_cta) {
      async_wait<0>();
      __syncthreads();
      acc.zero();
      tile_mma(As, Bs, acc);
      tile_sub(Ts, acc);
    }
  }
  async_wait<0>();
  __syncthreads();
  CHOL_TRACE(2);
  double r[4][4];
  if (!tile_cholesky(Ds, Lt, dinv, &fail, r)) {
    if (diag_cta && threadIdx.x == 0) sc->status = (double)((int64_t)j * NB + fail);
    return;
  }
  CHOL_TRACE(3);
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  if (diag_cta) {
    // G(j, j) lower, zeros above; G^T(j, j); log det over the first D pivots
    double* gtile = g + ((int64_t)j * NB) * Dp + (int64_t)j * NB;
    double* gttile = gt + ((int64_t)j * NB) * Dp + (int64_t)j * NB;
    double ld = 0.0;
#pragma unroll
    for (int aa = 0; aa < 4; ++aa)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const double v = ty >= tx ? r[aa][c] : 0.0;
        gtile[(int64_t)(4 * ty + aa) * Dp + 4 * tx + c] = v;
        gttile[(int64_t)(4 * tx + c) * Dp + 4 * ty + aa] = v;
        if (ty == tx && aa == c && (int64_t)j * NB + 4 * ty + aa < D) ld += log(v);
      }
    ld = block_sum(ld, red);
    if (threadIdx.x == 0) sc->logdet += 2.0 * ld;
    CHOL_TRACE(5);
    return;
  }
  // the factor, row-major (zero above the diagonal), into the diagonal tile's buffer: the DMMA solve reads it
#pragma unroll
  for (int aa = 0; aa < 4; ++aa)
#pragma unroll
    for (int c = 0; c < 4; ++c) Ds[(4 * ty + aa) * PITCH + 4 * tx + c] = ty >= tx ? r[aa][c] : 0.0;
  build_w16(Lt, dinv, w16);                  // ends with a block barrier: Ds and w16 visible to every warp
  tile_solve_mma(Ts, Ds, w16, rs, g + ((int64_t)i * NB) * Dp + (int64_t)j * NB,
                 gt + ((int64_t)j * NB) * Dp + (int64_t)i * NB, Dp);
  CHOL_TRACE(4);
  CHOL_TRACE(5);
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

// ---------------------------------------------------------------------------------------------
// Forward substitution with many right-hand sides, reduced to what the KL needs (ppi/policies.py:146-158):
//   out[0] = |L^-1 B|_F^2  (B lower triangular, D x D: tr(Sigma_old^-1 Sigma_new) with B = L_new),
//   out[1] = |L^-1 v|^2    ((mu_old - mu_new)^T Sigma_old^-1 (mu_old - mu_new)).
// CTA c owns the 64 columns c of B (CTA T: v as column 0 of a zero block) and walks down the block rows k:
//   Y_kc = L_kk^-1 (B_kc - sum_{j<k} L_kj Y_jc);
// the products run on the fp64 tensor cores (tile_mma wants the second operand as [column][contraction], so Y is
// kept TRANSPOSED in the workspace: Z_c[cc][i] = Y[i][64 c + cc]), the 64 x 64 substitution is warp-local (a warp
// owns 8 columns, 4 lanes per column split the dot product).  B lower triangular => Y_kc = 0 for k < c.
// Operand tiles of product j + 1 are in flight (cp.async) while product j multiplies.
// ---------------------------------------------------------------------------------------------
constexpr size_t kTrsmSmem = (size_t)(6 * NB * PITCH + NB) * sizeof(double);

__global__ void __launch_bounds__(kDenseThreads, 1)
trsm_sumsq_kernel(const double* __restrict__ Lp, int64_t Dp, int T, const double* __restrict__ B,
                  const double* __restrict__ v, int64_t D, double* __restrict__ Z, double* __restrict__ out) {
  extern __shared__ double smem[];
  double* Ab[2] = {smem, smem + 2 * NB * PITCH};
  double* Bb[2] = {smem + NB * PITCH, smem + 3 * NB * PITCH};
  double* Ts = smem + 4 * NB * PITCH;
  double* Ls = smem + 5 * NB * PITCH;
  double* invd = smem + 6 * NB * PITCH;
  __shared__ double red[32];
  const int c = blockIdx.x;
  const bool vec = c == T;
  if (vec ? v == nullptr : B == nullptr) return;
  double* Zc = Z + (int64_t)c * NB * Dp;
  const int k_first = vec ? 0 : c;
  const int tid = threadIdx.x, lane = tid & 31;
  double ss = 0.0;
  for (int k = k_first; k < T; ++k) {
    // right-hand side tile and the diagonal tile of L
    const int rows = (int)min((int64_t)NB, D - (int64_t)k * NB);
    if (!vec) {
      load_tile<true>(Ts, B + ((int64_t)k * NB) * D + (int64_t)c * NB, D, rows, (int)min((int64_t)NB, D - (int64_t)c * NB));
    } else {
      for (int e = tid; e < NB * NB; e += kDenseThreads) {
        const int r = e >> 6, cc = e & 63;
        Ts[r * PITCH + cc] = (cc == 0 && r < rows) ? v[(int64_t)k * NB + r] : 0.0;
      }
    }
    load_tile<false>(Ls, Lp + ((int64_t)k * NB) * Dp + (int64_t)k * NB, Dp);
    TileAcc acc;
    acc.zero();
    if (k > k_first) {
      load_tile_async(Ab[0], Lp + ((int64_t)k * NB) * Dp + (int64_t)k_first * NB, Dp);
      load_tile_async(Bb[0], Zc + (int64_t)k_first * NB, Dp);
      async_commit();
      for (int j = k_first; j < k; ++j) {
        const int cur = (j - k_first) & 1;
        if (j + 1 < k) {
          load_tile_async(Ab[cur ^ 1], Lp + ((int64_t)k * NB) * Dp + (int64_t)(j + 1) * NB, Dp);
          load_tile_async(Bb[cur ^ 1], Zc + (int64_t)(j + 1) * NB, Dp);
          async_commit();
          async_wait<1>();
        } else {
          async_wait<0>();
        }
        __syncthreads();
        tile_mma(Ab[cur], Bb[cur], acc);
        __syncthreads();
      }
    } else {
      __syncthreads();
    }
    tile_sub(Ts, acc);
    if (tid < NB) invd[tid] = 1.0 / Ls[tid * PITCH + tid];
    __syncthreads();
    // warp-local substitution: column cc = tid / 4, lane part q = tid % 4 sums over j = q, q + 4, ...
    {
      const int cc = tid >> 2, q = tid & 3;
      for (int i = 0; i < NB; ++i) {
        double sacc = 0.0;
        for (int j = q; j < i; j += 4) sacc = fma(Ls[i * PITCH + j], Ts[j * PITCH + cc], sacc);
        sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
        sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
        if (q == 0) {
          const double y = (Ts[i * PITCH + cc] - sacc) * invd[i];
          Ts[i * PITCH + cc] = y;
          ss = fma(y, y, ss);
        }
        __syncwarp();
      }
    }
    __syncthreads();
    // Z_c[cc][64 k + i] = Y_kc[i][cc]
    for (int e = tid; e < NB * NB; e += kDenseThreads) {
      const int cc = e >> 6, i = e & 63;
      Zc[(int64_t)cc * Dp + (int64_t)k * NB + i] = Ts[i * PITCH + cc];
    }
    __syncthreads();
  }
  ss = block_sum(ss, red);
  if (tid == 0) atomicAdd(out + (vec ? 1 : 0), ss);
  (void)lane;
}

constexpr size_t kTileSmem = 2 * NB * PITCH * sizeof(double);   // 69 632 B: opt-in above 48 KB
constexpr size_t kStepSmem = (size_t)(4 * NB * PITCH + 16 * kLtBlock + 256 + 4 * 16 * WP16 + 8 * 8 * WP16) * sizeof(double);

static int64_t pad64(int64_t D) { return (D + NB - 1) / NB * NB; }

static int set_dense_attrs() {
  static bool done = false;
  if (done) return 0;
  PPI_CUDA_TRY(cudaFuncSetAttribute(chol_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStepSmem));
  PPI_CUDA_TRY(cudaFuncSetAttribute(dgemm_nt_kernel<MODE_LG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTileSmem));
  PPI_CUDA_TRY(cudaFuncSetAttribute(dgemm_nt_kernel<MODE_LLT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTileSmem));
  PPI_CUDA_TRY(cudaFuncSetAttribute(trsm_sumsq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTrsmSmem));
  done = true;
  return 0;
}

// workspace: [scalars 1 KB][work Dp^2][G Dp^2][G^T Dp^2]
struct DenseWs {
  CholScalars* sc;
  double *work, *g, *gt;
};
static size_t dense_ws_bytes(int64_t D) {
  const int64_t Dp = pad64(D);
  return 1024 + (size_t)(3 * Dp * Dp) * sizeof(double) + 256;
}
static DenseWs dense_ws(void* workspace, int64_t D) {
  const int64_t Dp = pad64(D);
  char* p = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  DenseWs w;
  w.sc = (CholScalars*)p;
  w.work = (double*)(p + 1024);
  w.g = w.work + Dp * Dp;
  w.gt = w.g + Dp * Dp;
  return w;
}

// blocked factorisation of the padded matrix in ws.work -> ws.g / ws.gt, scalars in ws.sc: one launch per
// block column
static int chol_blocked(const DenseWs& ws, int64_t D, cudaStream_t st) {
  const int64_t Dp = pad64(D);
  const int T = (int)(Dp / NB);
  for (int j = 0; j < T; ++j) {
    const int nA = T - j, n = T - j - 1;
    const int nB = j > 0 ? n * (n + 1) / 2 : 0;
    chol_step_kernel<<<nA + nB, kDenseThreads, kStepSmem, st>>>(ws.work, Dp, j, T, D, ws.g, ws.gt, ws.sc);
  }
  return check_launch("blocked Cholesky kernels", T);
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

// tr(Sigma_old^-1 Sigma_new) and the Mahalanobis term of the KL from the two Cholesky factors:
// out[0] = |l^-1 b_lower|_F^2, out[1] = |l^-1 v|^2 (either operand may be NULL: its output is 0).
// workspace: ppi_dense_workspace_bytes(D) (padded copy of l + the transposed solution).
extern "C" int ppi_tri_solve_sumsq(const double* l, int64_t D, const double* b_lower, const double* v, double* out,
                                   void* workspace, size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(l && D > 0 && out && workspace && (b_lower || v), PPI_ERR_INVALID, "ppi_tri_solve_sumsq: bad arguments");
  PPI_REQUIRE(workspace_bytes >= dense_ws_bytes(D), PPI_ERR_WORKSPACE, "ppi_tri_solve_sumsq: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = set_dense_attrs()) return rc;
  const DenseWs ws = dense_ws(workspace, D);
  const int64_t Dp = pad64(D);
  const int T = (int)(Dp / NB);
  ProfileScope prof(PPI_PROF_DENSE_TAIL, st);
  PPI_CUDA_TRY(cudaMemsetAsync(out, 0, 2 * sizeof(double), st));
  pad_lower_kernel<<<(unsigned)min((int64_t)kNumSMs * 8, ceil_div(Dp * Dp, 256)), 256, 0, st>>>(l, D, Dp, ws.work);
  // Z = (T + 1) x 64 x Dp doubles: ws.g and the head of ws.gt (contiguous)
  trsm_sumsq_kernel<<<T + 1, kDenseThreads, kTrsmSmem, st>>>(ws.work, Dp, T, b_lower, v, D, ws.g, out);
  return check_launch("pad_lower_kernel / trsm_sumsq_kernel", 2);
}

// debug: the phase clocks of the last factorisation (64 steps x 2 CTAs x 8 slots), see g_chol_trace
extern "C" int ppi_debug_chol_trace(long long* host_out) {
  PPI_CUDA_TRY(cudaMemcpyFromSymbol(host_out, g_chol_trace, sizeof(long long) * 64 * 2 * 8));
  return 0;
}
