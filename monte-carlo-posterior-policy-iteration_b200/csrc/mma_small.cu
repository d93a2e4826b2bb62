// Warp-level tensor-core kernels for the small-matrix forms of the update path (H <= 32):
//   wmoments_matnorm_mma   acc += sum_b nw_b A_b diag(1/cov_out) A_b^T        ppi/policies.py:129-132
//   sample_matnorm_mma     x_b = clip(mean_fn + mean + L_in z_b diag(l_out))  ppi/policies.py:735-745
//
// These products are HBM-bound only if the H^2 d multiply-adds per sample cost (almost) nothing:
// at C3 shapes (H = 30, d = 28) they need ~100 TFLOP/s at HBM speed, more than the fp32 SIMT
// pipes deliver.  The matrices are far too small for a 128-row tcgen05 tile (a 30 x 30 output
// would waste > 90 % of it), so they run on warp-level mma.sync.m16n8k16 (fp16 in, fp32
// accumulate; measured 555 TFLOP/s on B200, tools/micro/mma_sync_rate.cu) with every operand in
// registers: one warp owns a 32 x 32 output, fragments are loaded straight from global memory in
// fragment order (every 32-byte sector fetched is fully used) and no shared-memory staging,
// swizzle or TMEM is involved.
//
// fp32 accuracy from fp16 tensor cores = the FP16X2 scheme of tc_gemm.cu: v 2^e = h1 + h2, three
// MMAs h1 g1 + h1 g2 + h2 g1, with the power-of-two scale chosen PER 32-COLUMN CHUNK by a warp max
// (the accumulator of a chunk is promoted to fp32 registers with an exact 2^-2e factor right after
// its 6 MMAs, so accumulation chains inside the tensor core stay 6 instructions long).
#include <cuda_fp16.h>

#include "common.cuh"

namespace ppi {

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// (h1, h2) halves of two scaled values packed as half2 registers (low half = first value)
__device__ __forceinline__ void split_pair(float v0, float v1, uint32_t& h1, uint32_t& h2) {
  const __half2 a = __floats2half2_rn(v0, v1);
  const float2 af = __half22float2(a);
  const __half2 b = __floats2half2_rn(v0 - af.x, v1 - af.y);
  h1 = *reinterpret_cast<const uint32_t*>(&a);
  h2 = *reinterpret_cast<const uint32_t*>(&b);
}

// predicated read-only loads (no branch: the compiler turns an `if` around a vector load into a
// divergent region)
__device__ __forceinline__ float2 ldg2_if(const float* p, bool pred) {
  float2 v;
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "setp.ne.b32 q, %3, 0;\n\t"
      "mov.f32 %0, 0f00000000;\n\t"
      "mov.f32 %1, 0f00000000;\n\t"
      "@q ld.global.nc.v2.f32 {%0, %1}, [%2];\n\t}"
      : "=f"(v.x), "=f"(v.y)
      : "l"(p), "r"((int)pred));
  return v;
}
__device__ __forceinline__ float ldg1_if(const float* p, bool pred) {
  float v;
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "setp.ne.b32 q, %2, 0;\n\t"
      "mov.f32 %0, 0f00000000;\n\t"
      "@q ld.global.nc.f32 %0, [%1];\n\t}"
      : "=f"(v)
      : "l"(p), "r"((int)pred));
  return v;
}

// single-instruction square root (MUFU.SQRT, relative error <= 2^-23, no slow path)
__device__ __forceinline__ float sqrt_fast(float v) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}

// 2^e with amax 2^e in [2^13, 2^14)
__device__ __forceinline__ float chunk_scale(float amax) {
  const int e = (int)((__float_as_uint(amax) >> 23) & 0xff) - 126;  // amax = m 2^e, m in [0.5, 1)
  int s = 14 - e;
  s = max(-60, min(60, s));   // inf / nan (exponent 255) clamp to 2^-60, zero never gets here
  return __uint_as_float((uint32_t)(s + 127) << 23);
}

constexpr int kMmaWarps = 8;  // warps per block

// ------------------------------------------------------------------------------------------
// acc[i,l] += sum_{b,j} nw_b inv_j (x_bij - s_ij)(x_blj - s_lj),   i <= l,   H <= 32.
// The contraction axis is the flattened (b, j) column index c = b d + j; a warp takes chunks of 32
// consecutive columns (any d, ragged across samples).  Per chunk and thread: rows {g, g+8, g+16,
// g+24} x columns {2t, 2t+1, 2t+8, 2t+9} (+16 for the second k-step), g = lane / 4, t = lane % 4:
// exactly the A (and, transposed, B) fragments of the 6 upper-triangular 16 x 8 output tiles.
// VEC2 (d even, x 8-byte aligned): the column pair (2t, 2t+1) is one 8-byte load.
// The loop body is branch-free: masked rows read a zero pad behind the shift table and their
// (predicated) load is skipped; zero-weight samples (possibly NaN rollouts) are never loaded.
// ------------------------------------------------------------------------------------------
template <bool VEC2>
__global__ void __launch_bounds__(kMmaWarps * 32)
wmoments_matnorm_mma_kernel(const float* __restrict__ x, const float* __restrict__ nw, int64_t K, int H,
                            int d, const float* __restrict__ shift, const float* __restrict__ inv_cov,
                            double* __restrict__ acc) {
  extern __shared__ float smem[];
  const int HD = H * d;
  const int HDp = (HD + 3) & ~3;           // the zero pad (d + 2 values) starts here
  const int padn = (d + 5) & ~3;
  float* sShift = smem;                    // HD values, then zeros
  float* sSqInv = sShift + HDp + padn;     // sqrt(inv_cov_out[j]), d values (+ pad)
  float* sRed = sSqInv + ((d + 3) & ~3);   // 32*32 block reduction buffer
  for (int i = threadIdx.x; i < HDp + padn; i += blockDim.x) sShift[i] = (i < HD && shift) ? shift[i] : 0.f;
  for (int i = threadIdx.x; i < ((d + 3) & ~3); i += blockDim.x)
    sSqInv[i] = i < d ? (inv_cov ? sqrtf(inv_cov[i]) : 1.f) : 0.f;
  for (int i = threadIdx.x; i < 32 * 32; i += blockDim.x) sRed[i] = 0.f;
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int64_t ncols = K * (int64_t)d;
  const int64_t nchunks = (ncols + 31) / 32;
  const int64_t wid = (int64_t)blockIdx.x * kMmaWarps + warp, nwarps = (int64_t)gridDim.x * kMmaWarps;
  const float inv_d = 1.0f / (float)d;
  const bool cols32 = ncols < (int64_t)0x7fffffff;  // 32-bit column arithmetic (no 64-bit division)

  // element offsets of this thread's 4 row groups inside a sample / inside the shift table
  int roff[4], soff[4];
  bool rok[4];
#pragma unroll
  for (int rg = 0; rg < 4; ++rg) {
    const int i = g + 8 * rg;
    rok[rg] = i < H;
    roff[rg] = rok[rg] ? i * d : 0;
    soff[rg] = rok[rg] ? i * d : HDp;  // masked rows: the zero pad behind the table
  }

  // tiles: (r,s) = (0,0) (0,1) (0,2) (0,3) (1,2) (1,3)
  float acc2[6][4];
#pragma unroll
  for (int q = 0; q < 6; ++q) acc2[q][0] = acc2[q][1] = acc2[q][2] = acc2[q][3] = 0.f;

  for (int64_t ch = wid; ch < nchunks; ch += nwarps) {
    const int64_t c0 = ch * 32;
    const int64_t b0 = cols32 ? (int64_t)((uint32_t)c0 / (uint32_t)d) : c0 / d;
    const int rel0 = (int)(c0 - b0 * d);
    const int ncol_here = (int)min((int64_t)32, ncols - c0);
    const float* xb = x + b0 * HD;  // offsets below fit 32 bits: <= (32 / d + 1) HD
    const float* wb = nw + b0;
    float dv[2][4][4];              // x - shift, [k-step][row group][4 columns]
    float fcol[2][4];               // sqrt(nw_b inv_j) per column
    float amax = 0.f;
#pragma unroll
    for (int st = 0; st < 2; ++st) {
#pragma unroll
      for (int cp = 0; cp < 2; ++cp) {  // column pair 2t, 2t+1 (+8)
        const int lc = 16 * st + 8 * cp + 2 * t;
        if (VEC2) {
          const int rel = rel0 + lc;
          const int bq = __float2int_rd(((float)rel + 0.5f) * inv_d);
          const int j = rel - bq * d;
          const float w = ldg1_if(wb + bq, lc < ncol_here);
          const float sw = sqrt_fast(w);
          const float2 si = *reinterpret_cast<const float2*>(sSqInv + j);
          fcol[st][2 * cp] = sw * si.x;
          fcol[st][2 * cp + 1] = sw * si.y;
          const int off = bq * HD + j;
          float m = 0.f;
#pragma unroll
          for (int rg = 0; rg < 4; ++rg) {
            const float2 xv = ldg2_if(xb + off + roff[rg], w != 0.f && rok[rg]);
            const float2 sv = *reinterpret_cast<const float2*>(sShift + soff[rg] + j);
            const float d0 = xv.x - sv.x, d1 = xv.y - sv.y;
            dv[st][rg][2 * cp] = d0;
            dv[st][rg][2 * cp + 1] = d1;
            m = fmaxf(m, fmaxf(fabsf(d0) * si.x, fabsf(d1) * si.y));
          }
          amax = fmaxf(amax, m * sw);
        } else {
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int rel = rel0 + lc + u;
            const int bq = __float2int_rd(((float)rel + 0.5f) * inv_d);
            const int j = rel - bq * d;
            const float w = ldg1_if(wb + bq, (lc + u) < ncol_here);
            const float f = sqrt_fast(w) * sSqInv[j];
            fcol[st][2 * cp + u] = f;
            const int off = bq * HD + j;
            float m = 0.f;
#pragma unroll
            for (int rg = 0; rg < 4; ++rg) {
              const float xv = ldg1_if(xb + off + roff[rg], w != 0.f && rok[rg]);
              const float d0 = xv - sShift[soff[rg] + j];
              dv[st][rg][2 * cp + u] = d0;
              m = fmaxf(m, fabsf(d0));
            }
            amax = fmaxf(amax, m * f);
          }
        }
      }
    }
    amax = warp_max(amax);
    if (amax == 0.f) continue;  // warp-uniform: nothing but zero-weight samples / padding
    const float sc = chunk_scale(amax);
    const float isc2 = (1.f / sc) * (1.f / sc);

    float a1[6][4];
#pragma unroll
    for (int q = 0; q < 6; ++q) a1[q][0] = a1[q][1] = a1[q][2] = a1[q][3] = 0.f;
#pragma unroll
    for (int st = 0; st < 2; ++st) {
      float fs[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) fs[c] = fcol[st][c] * sc;
      uint32_t h1[4][2], h2[4][2];  // [row group][column pair]
#pragma unroll
      for (int rg = 0; rg < 4; ++rg)
#pragma unroll
        for (int cp = 0; cp < 2; ++cp)
          split_pair(dv[st][rg][2 * cp] * fs[2 * cp], dv[st][rg][2 * cp + 1] * fs[2 * cp + 1], h1[rg][cp],
                     h2[rg][cp]);
      // tile q = (r, s): A rows 16r + {g, g+8} -> row groups 2r, 2r+1; B rows 8s + g -> row group s
#define PPI_TILE(q, r, s)                                                                              \
  mma16816(a1[q], h1[2 * r][0], h1[2 * r + 1][0], h1[2 * r][1], h1[2 * r + 1][1], h1[s][0], h1[s][1]); \
  mma16816(a1[q], h1[2 * r][0], h1[2 * r + 1][0], h1[2 * r][1], h1[2 * r + 1][1], h2[s][0], h2[s][1]); \
  mma16816(a1[q], h2[2 * r][0], h2[2 * r + 1][0], h2[2 * r][1], h2[2 * r + 1][1], h1[s][0], h1[s][1]);
      PPI_TILE(0, 0, 0) PPI_TILE(1, 0, 1) PPI_TILE(2, 0, 2) PPI_TILE(3, 0, 3) PPI_TILE(4, 1, 2) PPI_TILE(5, 1, 3)
#undef PPI_TILE
    }
#pragma unroll
    for (int q = 0; q < 6; ++q)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc2[q][e] = fmaf(a1[q][e], isc2, acc2[q][e]);
  }

  // block reduction in shared memory (fp32 atomics on <= 8 warps), then fp64 atomics to global
  const int tr[6] = {0, 0, 0, 0, 1, 1}, ts[6] = {0, 1, 2, 3, 2, 3};
#pragma unroll
  for (int q = 0; q < 6; ++q) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = 16 * tr[q] + g + 8 * (e >> 1), l = 8 * ts[q] + 2 * t + (e & 1);
      atomicAdd(&sRed[i * 32 + l], acc2[q][e]);
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < 32 * 32; e += blockDim.x) {
    const int i = e >> 5, l = e & 31;
    if (i < H && l < H && l >= i) {
      const float s = sRed[e];
      if (s != 0.f) atomicAdd(acc + (int64_t)i * H + l, (double)s);
    }
  }
}


// ------------------------------------------------------------------------------------------
// x_b = clip(mean_fn + mean + L_in z_b diag(l_out)),  H <= 32, d <= 32, one warp per sample.
// (L_in z_b) runs as M = 32 rows x N = d columns x K = 32 contraction rows of mma.sync; the
// diagonal output factor is applied to the fp32 result in the epilogue, so in Philox mode the B
// operand is the raw normal, an exact single fp16 (2 products L1 z + L2 z); injected normals are
// split into an fp16 pair under a per-sample power-of-two scale (3 products).  The warp first
// builds the sample's z tile in shared memory ([row][column] halves, row stride 40 halves = 80 B:
// 16-byte aligned, conflict-free rows for ldmatrix); the B fragments, which pack two contraction
// rows per register, come out of ldmatrix.x4.trans, one instruction per 8 output columns; the L_in
// fragments (3 non-zero 16 x 16 blocks of the lower triangle, pair-split under a
// block-wide scale) stay in registers for the whole kernel.
// ------------------------------------------------------------------------------------------
// B fragments of one 8-column tile for both k-steps: four 8 x 8 (k x n) blocks, transposed on the fly
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                  uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

constexpr int kTileStride = 40;              // halves per tile row (contraction row i' of z)
constexpr int kTileHalves = 32 * kTileStride;

template <bool PHILOX, bool VEC2>
__global__ void __launch_bounds__(kMmaWarps * 32)
sample_matnorm_mma_kernel(const float* __restrict__ z, uint64_t seed, uint64_t draw, uint64_t row_offset,
                          int64_t K, int H, int d, const float* __restrict__ mean_fn,
                          const float* __restrict__ mean, const float* __restrict__ l_in,
                          const float* __restrict__ l_out, const float* __restrict__ lower,
                          const float* __restrict__ upper, float* __restrict__ x) {
  extern __shared__ float smem[];
  const int HD = H * d;
  float* sBias = smem;                       // HD (+ pad to even)
  float* sFo = sBias + ((HD + 3) & ~3);      // 32: l_out[j][j]
  float* sLo = sFo + 32;                     // 32
  float* sHi = sLo + 32;                     // 32
  float* sMax = sHi + 32;                    // 32 (block max of |L_in|)
  __half* sTiles = reinterpret_cast<__half*>(sMax + 32);  // per warp: T1 (and T2 when !PHILOX)
  constexpr int kTilesPerWarp = PHILOX ? 1 : 2;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;

  float lmax = 0.f;
  for (int i = threadIdx.x; i < HD; i += blockDim.x) sBias[i] = mean_fn[i % d] + mean[i];
  for (int i = threadIdx.x; i < 32; i += blockDim.x) {
    sFo[i] = i < d ? l_out[i * d + i] : 0.f;
    sLo[i] = (lower && i < d) ? lower[i] : -INFINITY;
    sHi[i] = (upper && i < d) ? upper[i] : INFINITY;
  }
  for (int i = threadIdx.x; i < H * H; i += blockDim.x) lmax = fmaxf(lmax, fabsf(l_in[i]));
  for (int i = threadIdx.x; i < kMmaWarps * kTilesPerWarp * kTileHalves; i += blockDim.x)
    sTiles[i] = __float2half(0.f);
  lmax = warp_max(lmax);
  if (lane == 0) sMax[warp] = lmax;
  __syncthreads();
  lmax = 0.f;
#pragma unroll
  for (int w = 0; w < kMmaWarps; ++w) lmax = fmaxf(lmax, sMax[w]);
  const float sL = lmax > 0.f ? chunk_scale(lmax) : 1.f;

  // A fragments of L_in 2^eL: blocks (m,k) = (0,0), (1,0), (1,1)
  uint32_t A1[3][4], A2[3][4];
  {
    const int bm[3] = {0, 1, 1}, bk[3] = {0, 0, 1};
#pragma unroll
    for (int q = 0; q < 3; ++q) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {  // a0: (g, 2t) a1: (g+8, 2t) a2: (g, 2t+8) a3: (g+8, 2t+8)
        const int i = 16 * bm[q] + g + 8 * (e & 1), k = 16 * bk[q] + 2 * t + 8 * (e >> 1);
        const float v0 = (i < H && k < H && k <= i) ? l_in[i * H + k] * sL : 0.f;
        const float v1 = (i < H && k + 1 < H && k + 1 <= i) ? l_in[i * H + k + 1] * sL : 0.f;
        split_pair(v0, v1, A1[q][e], A2[q][e]);
      }
    }
  }

  __half* T1 = sTiles + (size_t)warp * kTilesPerWarp * kTileHalves;
  __half* T2 = T1 + kTileHalves;  // only when !PHILOX
  // ldmatrix row address of this lane: block lane / 8 (contraction rows 8 blk ..), row lane % 8
  const uint32_t ld1 = (uint32_t)__cvta_generic_to_shared(T1 + lane * kTileStride);
  const uint32_t ld2 = (uint32_t)__cvta_generic_to_shared(T2 + lane * kTileStride);
  const PhiloxKeys keys = philox_keys(seed);
  const bool quad_rows = (d & 3) == 0;
  const int nt = (d + 7) >> 3;
  const float inv_d = 1.0f / (float)d;
  const int64_t wid = (int64_t)blockIdx.x * kMmaWarps + warp, nwarps = (int64_t)gridDim.x * kMmaWarps;

  for (int64_t b = wid; b < K; b += nwarps) {
    float inv_scale = 1.f / sL;
    __syncwarp();  // the previous sample's fragment reads are done
    if (PHILOX) {
      const uint64_t e0 = (row_offset + (uint64_t)b) * (uint64_t)HD;
      const uint64_t q0 = e0 >> 2;
      const int lead = (int)(e0 & 3);          // elements of the first quad that belong to the previous sample
      const int nq = (HD + lead + 3) >> 2;
      for (int qi = lane; qi < nq; qi += 32) {
        float v[4];
        philox_normal4(keys, draw, q0 + (uint64_t)qi, v);
        const int el0 = 4 * qi - lead;
        int ip = __float2int_rd(((float)max(el0, 0) + 0.5f) * inv_d);
        int j = max(el0, 0) - ip * d;
        if (quad_rows) {  // d % 4 == 0: a quad never straddles a row and is 8-byte aligned in the tile
          const __half2 p0 = __floats2half2_rn(v[0], v[1]), p1 = __floats2half2_rn(v[2], v[3]);  // exact
          *reinterpret_cast<uint2*>(T1 + ip * kTileStride + j) =
              make_uint2(*reinterpret_cast<const uint32_t*>(&p0), *reinterpret_cast<const uint32_t*>(&p1));
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int el = el0 + k;
            if (el >= 0 && el < HD) {
              T1[ip * kTileStride + j] = __float2half_rn(v[k]);  // exact: Philox normals are half values
              if (++j == d) { j = 0; ++ip; }
            }
          }
        }
      }
    } else {
      const float* zb = z + b * HD;
      float v[32];
      float amax = 0.f;
#pragma unroll
      for (int it = 0; it < 32; ++it) {
        const int el = it * 32 + lane;
        v[it] = el < HD ? __ldg(zb + el) : 0.f;
        amax = fmaxf(amax, fabsf(v[it]));
      }
      amax = warp_max(amax);
      const float sz = amax > 0.f ? chunk_scale(amax) : 1.f;
      inv_scale *= 1.f / sz;
#pragma unroll
      for (int it = 0; it < 32; ++it) {
        const int el = it * 32 + lane;
        if (el < HD) {
          const int ip = __float2int_rd(((float)el + 0.5f) * inv_d);
          const int j = el - ip * d;
          const float vs = v[it] * sz;
          const __half h1 = __float2half_rn(vs);
          T1[ip * kTileStride + j] = h1;
          T2[ip * kTileStride + j] = __float2half_rn(vs - __half2float(h1));
        }
      }
    }
    __syncwarp();

    float acc[2][4][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int s = 0; s < 4; ++s) acc[m][s][0] = acc[m][s][1] = acc[m][s][2] = acc[m][s][3] = 0.f;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      if (s < nt) {  // warp-uniform
        uint32_t b00, b01, b10, b11;  // rows (2t, 2t+1) + {0, 8, 16, 24} of column 8s + g
        ldmatrix_x4_trans(ld1 + 16u * s, b00, b01, b10, b11);
        mma16816(acc[0][s], A1[0][0], A1[0][1], A1[0][2], A1[0][3], b00, b01);
        mma16816(acc[0][s], A2[0][0], A2[0][1], A2[0][2], A2[0][3], b00, b01);
        mma16816(acc[1][s], A1[1][0], A1[1][1], A1[1][2], A1[1][3], b00, b01);
        mma16816(acc[1][s], A2[1][0], A2[1][1], A2[1][2], A2[1][3], b00, b01);
        mma16816(acc[1][s], A1[2][0], A1[2][1], A1[2][2], A1[2][3], b10, b11);
        mma16816(acc[1][s], A2[2][0], A2[2][1], A2[2][2], A2[2][3], b10, b11);
        if (!PHILOX) {
          uint32_t c00, c01, c10, c11;
          ldmatrix_x4_trans(ld2 + 16u * s, c00, c01, c10, c11);
          mma16816(acc[0][s], A1[0][0], A1[0][1], A1[0][2], A1[0][3], c00, c01);
          mma16816(acc[1][s], A1[1][0], A1[1][1], A1[1][2], A1[1][3], c00, c01);
          mma16816(acc[1][s], A1[2][0], A1[2][1], A1[2][2], A1[2][3], c10, c11);
        }
      }
    }

    // epilogue: D fragment (rows 16m + g (+8), columns 8s + 2t (+1)) -> bias, factor, clip, store
    float* xb = x + b * HD;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const int j0 = 8 * s + 2 * t;
      if (s < nt && j0 < d) {
        const bool two = j0 + 1 < d;
        const float f0 = sFo[j0] * inv_scale, f1 = sFo[j0 + 1] * inv_scale;
        const float lo0 = sLo[j0], lo1 = sLo[j0 + 1], hi0 = sHi[j0], hi1 = sHi[j0 + 1];
#pragma unroll
        for (int m = 0; m < 2; ++m) {
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const int i = 16 * m + g + 8 * hh;
            if (i < H) {
              const int o = i * d + j0;
              const float v0 = clipf(fmaf(acc[m][s][2 * hh], f0, sBias[o]), lo0, hi0);
              if (VEC2) {
                const float v1 = clipf(fmaf(acc[m][s][2 * hh + 1], f1, sBias[o + 1]), lo1, hi1);
                *reinterpret_cast<float2*>(xb + o) = make_float2(v0, v1);
              } else {
                xb[o] = v0;
                if (two) xb[o + 1] = clipf(fmaf(acc[m][s][2 * hh + 1], f1, sBias[o + 1]), lo1, hi1);
              }
            }
          }
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// Feature expansion  y_b = clip([mean_fn + Phi w_b, Phi' w_b])  for d = 2, F <= 32
// (ppi/policies.py:1009-1016: actions of an RBF / RFF feature policy from its sampled weights).
// One (T x F) x (F x 2K) product whose output (K, T, dy) is the only large HBM stream: 21 multiply-
// adds per output float on the SIMT pipes cost as much as writing it, so the product runs on
// mma.sync with the FP16X2 split instead.  A block owns 256 time steps: each warp keeps the (h1, h2)
// fragments of its 32 rows of Phi (and Phi') in registers for the whole kernel and walks over groups
// of 32 samples whose weights are split once per group (power-of-two scale from the group maximum)
// into a shared tile [column n = (sample, channel)][feature], 40 halves apart so that the B-fragment
// loads of a warp hit 32 different banks.  With n = 2 s + c the accumulator pair of a thread is the
// two channels of one sample at one time step: value and derivative leave as one 16-byte store,
// eight consecutive time steps of a sample forming one 128-byte line.
// ------------------------------------------------------------------------------------------
constexpr int kFxS = 32;       // samples per group
constexpr int kFxStride = 40;  // halves between columns of the split-weight tile
constexpr int kFxTile = 256;   // time steps per block

template <bool DERIV>
__global__ void __launch_bounds__(256)
features_expand_mma_kernel(const float* __restrict__ w, int64_t K, int F, int T,
                           const float* __restrict__ mean_fn, const float* __restrict__ feat,
                           const float* __restrict__ dfeat, const float* __restrict__ lower,
                           const float* __restrict__ upper, float* __restrict__ y) {
  __shared__ __align__(16) __half sW1[2 * kFxS * kFxStride];
  __shared__ __align__(16) __half sW2[2 * kFxS * kFxStride];
  __shared__ float sRed[16];
  constexpr int DY = DERIV ? 4 : 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
  const int tb = blockIdx.y * kFxTile;
  const int t0 = tb + warp * 32;

  // block-wide scales of the Phi / Phi' tiles
  float pmax = 0.f, dmax = 0.f;
  {
    const int nt = min(kFxTile, T - tb);
    for (int i = threadIdx.x; i < nt * F; i += blockDim.x) {
      pmax = fmaxf(pmax, fabsf(feat[(int64_t)tb * F + i]));
      if (DERIV) dmax = fmaxf(dmax, fabsf(dfeat[(int64_t)tb * F + i]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      pmax = fmaxf(pmax, __shfl_xor_sync(0xffffffffu, pmax, o));
      dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    }
    if (lane == 0) { sRed[warp] = pmax; sRed[8 + warp] = dmax; }
    for (int i = threadIdx.x; i < 2 * kFxS * kFxStride; i += blockDim.x) {
      sW1[i] = __float2half(0.f);
      sW2[i] = __float2half(0.f);
    }
    __syncthreads();
    pmax = dmax = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { pmax = fmaxf(pmax, sRed[i]); dmax = fmaxf(dmax, sRed[8 + i]); }
    __syncthreads();
  }
  const float psc = pmax > 0.f ? chunk_scale(pmax) : 1.f;
  const float dsc = dmax > 0.f ? chunk_scale(dmax) : 1.f;

  // A fragments: [row tile][k-step][register], (h1, h2) of Phi and Phi'
  uint32_t a1[2][2][4], a2[2][2][4], e1[2][2][4], e2[2][2][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int ks = 0; ks < 2; ++ks)
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int t = t0 + 16 * i + g + 8 * (r & 1);
        const int f = 16 * ks + 2 * q + 8 * (r >> 1);
        const bool ok0 = t < T && f < F, ok1 = t < T && f + 1 < F;
        const int64_t o = (int64_t)t * F + f;
        split_pair(ok0 ? feat[o] * psc : 0.f, ok1 ? feat[o + 1] * psc : 0.f, a1[i][ks][r], a2[i][ks][r]);
        if (DERIV)
          split_pair(ok0 ? dfeat[o] * dsc : 0.f, ok1 ? dfeat[o + 1] * dsc : 0.f, e1[i][ks][r], e2[i][ks][r]);
      }

  float lo[DY], hi[DY], mf[2];
#pragma unroll
  for (int c = 0; c < DY; ++c) {
    lo[c] = lower ? lower[c] : -INFINITY;
    hi[c] = upper ? upper[c] : INFINITY;
  }
  mf[0] = mean_fn[0];
  mf[1] = mean_fn[1];
  uint32_t tmask = 0;   // bit 2 i + hh: row t0 + 16 i + g + 8 hh exists
#pragma unroll
  for (int j = 0; j < 4; ++j) tmask |= (t0 + 16 * (j >> 1) + g + 8 * (j & 1) < T ? 1u : 0u) << j;

  const int FD = 2 * F;
  const int per_group = kFxS * FD;             // <= 2048 values: <= 8 per thread
  const int64_t n_groups = ceil_div(K, (int64_t)kFxS);
  float wv[8];
  auto fetch = [&](int64_t grp) {
    const int64_t base = grp * per_group;
    const int64_t lim = K * FD;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int e = threadIdx.x + 256 * j;
      wv[j] = ldg1_if(w + base + e, e < per_group && base + e < lim);
    }
  };
  if ((int64_t)blockIdx.x < n_groups) fetch(blockIdx.x);
  for (int64_t grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {
    const int64_t b0 = grp * kFxS;
    const int ns = (int)min((int64_t)kFxS, K - b0);
    float amax = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(wv[j]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (lane == 0) sRed[warp] = amax;
    __syncthreads();   // previous group's B loads are done as well
    amax = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) amax = fmaxf(amax, sRed[i]);
    const float wsc = amax > 0.f ? chunk_scale(amax) : 1.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int e = threadIdx.x + 256 * j;
      if (e < per_group) {
        const int s = e / FD, r = e - s * FD;       // r = 2 f + c
        const float v = wv[j] * wsc;
        const __half h1 = __float2half_rn(v);
        const __half h2 = __float2half_rn(v - __half2float(h1));
        const int o = (2 * s + (r & 1)) * kFxStride + (r >> 1);
        sW1[o] = h1;
        sW2[o] = h2;
      }
    }
    const int64_t nxt = grp + gridDim.x;
    if (nxt < n_groups) fetch(nxt);
    __syncthreads();
    const float py = 1.f / (psc * wsc), dy = 1.f / (dsc * wsc);
    const bool full = ns == kFxS;
    float* yb = y + ((b0 + q) * (int64_t)T + t0 + g) * DY;   // sample b0 + 4 nt + q, row t0 + g
#pragma unroll 2
    for (int nt = 0; nt < 8; ++nt, yb += 4 * (int64_t)T * DY) {
      const int col = (nt * 8 + g) * kFxStride + 2 * q;
      uint32_t b1[2][2], b2[2][2];
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          b1[ks][r] = *reinterpret_cast<const uint32_t*>(sW1 + col + 16 * ks + 8 * r);
          b2[ks][r] = *reinterpret_cast<const uint32_t*>(sW2 + col + 16 * ks + 8 * r);
        }
      // four independent accumulation chains (2 row tiles x {Phi, Phi'}), issued round-robin
      float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      float dac[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
#pragma unroll
        for (int p = 0; p < 3; ++p)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const uint32_t* fa = p == 2 ? a2[i][ks] : a1[i][ks];
            const uint32_t* fb = p == 1 ? b2[ks] : b1[ks];
            mma16816(acc[i], fa[0], fa[1], fa[2], fa[3], fb[0], fb[1]);
            if (DERIV) {
              const uint32_t* fe = p == 2 ? e2[i][ks] : e1[i][ks];
              mma16816(dac[i], fe[0], fe[1], fe[2], fe[3], fb[0], fb[1]);
            }
          }
      if (full || nt * 4 + q < ns) {
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            if (tmask >> (2 * i + hh) & 1) {
              const float v0 = clipf(fmaf(acc[i][2 * hh], py, mf[0]), lo[0], hi[0]);
              const float v1 = clipf(fmaf(acc[i][2 * hh + 1], py, mf[1]), lo[1], hi[1]);
              if (DERIV) {
                const float v2 = clipf(dac[i][2 * hh] * dy, lo[2], hi[2]);
                const float v3 = clipf(dac[i][2 * hh + 1] * dy, lo[3], hi[3]);
                *reinterpret_cast<float4*>(yb + (16 * i + 8 * hh) * 4) = make_float4(v0, v1, v2, v3);
              } else {
                *reinterpret_cast<float2*>(yb + (16 * i + 8 * hh) * 2) = make_float2(v0, v1);
              }
            }
          }
      }
    }
  }
}

}  // namespace ppi

using namespace ppi;

// called from moments.cu for H <= 32
int ppi_wmoments_matnorm_mma(const float* x, const float* nw, int64_t K, int H, int d, const float* shift,
                             const float* inv_cov_out, double* acc, cudaStream_t st) {
  const size_t smem = ((size_t)H * d + 2 * (size_t)d + 24 + 32 * 32) * sizeof(float);
  PPI_REQUIRE(smem <= 200 * 1024, PPI_ERR_UNSUPPORTED,
              "ppi_wmoments_matnorm: H=%d d=%d does not fit shared memory", H, d);
  const bool vec2 = (d % 2 == 0) && ((reinterpret_cast<uintptr_t>(x) & 7) == 0);
  if (smem > 48 * 1024) {
    PPI_CUDA_TRY(cudaFuncSetAttribute(wmoments_matnorm_mma_kernel<true>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PPI_CUDA_TRY(cudaFuncSetAttribute(wmoments_matnorm_mma_kernel<false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  const int64_t nchunks = (K * (int64_t)d + 31) / 32;
  int64_t grid = ceil_div(nchunks, (int64_t)kMmaWarps * 4);  // >= 4 chunks per warp
  if (grid > (int64_t)kNumSMs * 2) grid = (int64_t)kNumSMs * 2;
  if (grid < 1) grid = 1;
  if (vec2)
    wmoments_matnorm_mma_kernel<true><<<(unsigned)grid, kMmaWarps * 32, smem, st>>>(x, nw, K, H, d, shift,
                                                                                   inv_cov_out, acc);
  else
    wmoments_matnorm_mma_kernel<false><<<(unsigned)grid, kMmaWarps * 32, smem, st>>>(x, nw, K, H, d, shift,
                                                                                    inv_cov_out, acc);
  return check_launch("wmoments_matnorm_mma_kernel");
}

// called from sample.cu for H <= 32, d <= 32, diagonal l_out
int ppi_sample_matnorm_mma(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset, int64_t K,
                           int H, int d, const float* mean_fn, const float* mean, const float* l_in,
                           const float* l_out, const float* lower, const float* upper, float* x,
                           cudaStream_t st) {
  const size_t tiles = (size_t)kMmaWarps * (z ? 2 : 1) * kTileHalves * sizeof(__half);
  const size_t smem = ((((size_t)H * d + 3) & ~(size_t)3) + 4 * 32) * sizeof(float) + tiles;
  const bool vec2 = (d % 2 == 0) && ((reinterpret_cast<uintptr_t>(x) & 7) == 0);
  int64_t grid = ceil_div(K, (int64_t)kMmaWarps * 2);  // >= 2 samples per warp
  if (grid > (int64_t)kNumSMs * 2) grid = (int64_t)kNumSMs * 2;
  if (grid < 1) grid = 1;
#define PPI_LAUNCH_SMM(P, V)                                                                            \
  do {                                                                                                  \
    if (smem > 48 * 1024)                                                                               \
      PPI_CUDA_TRY(cudaFuncSetAttribute(sample_matnorm_mma_kernel<P, V>,                                \
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    sample_matnorm_mma_kernel<P, V><<<(unsigned)grid, kMmaWarps * 32, smem, st>>>(                      \
        z, seed, draw_id, row_offset, K, H, d, mean_fn, mean, l_in, l_out, lower, upper, x);            \
  } while (0)
  if (z == nullptr) { if (vec2) PPI_LAUNCH_SMM(true, true); else PPI_LAUNCH_SMM(true, false); }
  else              { if (vec2) PPI_LAUNCH_SMM(false, true); else PPI_LAUNCH_SMM(false, false); }
#undef PPI_LAUNCH_SMM
  return check_launch("sample_matnorm_mma_kernel");
}

// called from sample.cu: feature policies with d = 2, F <= 32 (weights w already sampled)
int ppi_features_expand_mma(const float* w, int64_t K, int F, int T, const float* mean_fn,
                            const float* feat, const float* dfeat, const float* lower,
                            const float* upper, float* y, cudaStream_t st) {
  const int ty = (int)ceil_div(T, kFxTile);
  int64_t gx = ceil_div((int64_t)kNumSMs * 2, (int64_t)ty);
  const int64_t n_groups = ceil_div(K, (int64_t)kFxS);
  if (gx > n_groups) gx = n_groups;
  if (gx < 1) gx = 1;
  dim3 grid((unsigned)gx, (unsigned)ty);
  if (dfeat)
    features_expand_mma_kernel<true><<<grid, 256, 0, st>>>(w, K, F, T, mean_fn, feat, dfeat, lower, upper, y);
  else
    features_expand_mma_kernel<false><<<grid, 256, 0, st>>>(w, K, F, T, mean_fn, feat, dfeat, lower, upper, y);
  return check_launch("features_expand_mma_kernel");
}
