// tcgen05 / TMA kernels of the PPI update path (sm_100a).
//
//   ppi_wmoments_dense   sum_k nw_k (x_k - s)(x_k - s)^T          ppi/policies.py:89-92
//   ppi_sample_gaussian  x = mu + z L^T  (D > 128)                ppi/policies.py:247-249,
//                                                                  ppi/samplers.py:38-39
//
// Both are the same tensor-core contraction  C[m,n] = sum_k A[m,k] B[n,k]  with both operands
// stored K-major (contraction index contiguous), which is the canonical tcgen05 shared-memory
// layout (128-byte swizzle, 8-row atoms).
//
// Precision.  tcgen05 has no fp32-input MMA; kind::tf32 keeps 10 mantissa bits.  To meet the
// 1e-5 parity bar every operand is split on the fly by a pre-pass into hi = tf32(v) and
// lo = tf32(v - hi) and each k-step issues three MMAs (hi*hi + hi*lo + lo*hi), which leaves a
// relative product error of ~2^-22.  The fp32 TMEM accumulator is drained into registers every
// `chunk_kb` k-blocks (two 256-column TMEM buffers ping-pong), so no accumulation chain inside
// the tensor core is longer than 3 * 4 * chunk_kb MMAs; across chunks the sum is carried in
// fp32 registers with round-to-nearest adds.
//
// CTA = 10 warps: warp 0 TMA producer (one lane), warp 1 MMA issuer (one lane) + TMEM
// allocation, warps 2..9 drain TMEM (each warp owns its hardware lane quarter and one 128-column
// half).  One output tile of 128 x 256 per CTA.
#include <cuda.h>

#include "common.cuh"

namespace ppi {

constexpr int BM = 128;        // UMMA M (TMEM lanes)
constexpr int BN = 256;        // UMMA N (TMEM columns per accumulator)
constexpr int BK = 32;         // fp32 elements per k-block = 128 bytes = one swizzle row
constexpr int UMMA_K = 8;      // tf32
constexpr int kStages = 2;
constexpr int kGemmThreads = 320;
constexpr uint32_t kABytes = BM * BK * 4;   // 16 KB
constexpr uint32_t kBBytes = BN * BK * 4;   // 32 KB
constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytes;  // Ah, Al, Bh, Bl = 96 KB
constexpr uint32_t kGemmSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;

enum { GEMM_SYRK = 0, GEMM_SAMPLE = 1 };

struct GemmParams {
  int mode;
  int num_kb;        // total k-blocks of the contraction
  int chunk_kb;      // k-blocks per TMEM accumulation chunk
  int kb_per_split;  // SYRK: k-blocks handled by one CTA
  int splits;        // SYRK: CTAs per output tile
  int tiles_m, tiles_n;
  int64_t M, N;      // valid output extent (rows, cols)
  float* out;        // SYRK: partials [splits][ld][ld]; SAMPLE: x [M][ld]
  int64_t ld;
  int64_t split_stride;
  const float* bias; // SAMPLE: mu[n]
  int accumulate;    // SYRK: partial += acc
  int a_exact;       // SAMPLE: A (the Philox normals) is exactly TF32: no A_lo operand, 2 MMAs per step
};

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_wait_ld() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128-byte swizzle, rows 128 B apart, 8-row atoms 1024 B apart (cute::UMMA::SmemDescriptor:
// start[0,14) LBO[16,30) SBO[32,46) version[46,48)=1 layout[61,64)=2).
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// kind::tf32, fp32 accumulate, A and B K-major (cute::UMMA::InstrDescriptor).
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) |
                            ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ uint32_t tf32_rna(float v) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
  return u;
}
__device__ __forceinline__ void split_tf32(float v, float& hi, float& lo) {
  hi = __uint_as_float(tf32_rna(v));
  lo = __uint_as_float(tf32_rna(v - hi));
}

// ---------------------------------------------------------------------------------------------
// The GEMM kernel
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kGemmThreads, 1)
split_tf32_gemm_kernel(const __grid_constant__ CUtensorMap map_ah,
                       const __grid_constant__ CUtensorMap map_al,
                       const __grid_constant__ CUtensorMap map_bh,
                       const __grid_constant__ CUtensorMap map_bl, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw);
  const uint32_t bar_base = base + kStages * kStageBytes;
  // barriers: full[kStages], empty[kStages], tmem_full[2], tmem_empty[2], then the TMEM pointer
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tfull_bar = [&](int b) { return bar_base + 8u * (2 * kStages + b); };
  auto tempty_bar = [&](int b) { return bar_base + 8u * (2 * kStages + 2 + b); };
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem + kStages * kStageBytes + 8 * (2 * kStages + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- tile / k-range of this CTA ----
  int tm, tn, kb_begin, kb_end, split = 0;
  if (p.mode == GEMM_SYRK) {
    int t = blockIdx.x / p.splits;
    split = blockIdx.x % p.splits;
    tm = 0;
    for (;; ++tm) {  // upper-triangular tiles: tn >= tn_min(tm)
      const int tn_min = (tm * BM) / BN;
      const int cnt = p.tiles_n - tn_min;
      if (t < cnt) { tn = tn_min + t; break; }
      t -= cnt;
    }
    kb_begin = split * p.kb_per_split;
    kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
  } else {
    tm = blockIdx.x / p.tiles_n;
    tn = blockIdx.x % p.tiles_n;
    kb_begin = 0;
    kb_end = min(p.num_kb, ((tn + 1) * BN + BK - 1) / BK);  // L is lower triangular
  }
  const int nkb = max(0, kb_end - kb_begin);
  const int nchunks = (nkb + p.chunk_kb - 1) / p.chunk_kb;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(tfull_bar(b), 1); mbar_init(tempty_bar(b), 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32((const void*)tmem_ptr_smem)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    if (lane == 0 && nkb > 0) {
      // ================= TMA producer =================
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_ah) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_al) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_bh) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_bl) : "memory");
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t sb = base + stage * kStageBytes;
        mbar_expect_tx(full_bar(stage), p.a_exact ? kStageBytes - kABytes : kStageBytes);
        tma_load_2d(sb, &map_ah, full_bar(stage), kb * BK, tm * BM);
        if (!p.a_exact) tma_load_2d(sb + kABytes, &map_al, full_bar(stage), kb * BK, tm * BM);
        tma_load_2d(sb + 2 * kABytes, &map_bh, full_bar(stage), kb * BK, tn * BN);
        tma_load_2d(sb + 2 * kABytes + kBBytes, &map_bl, full_bar(stage), kb * BK, tn * BN);
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && nkb > 0) {
      // ================= MMA issuer =================
      int stage = 0;
      uint32_t phase = 0;
      int kb = 0;
      for (int c = 0; c < nchunks; ++c) {
        const int buf = c & 1;
        mbar_wait(tempty_bar(buf), (((uint32_t)c >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)buf * BN;
        const int kb_chunk_end = min(nkb, kb + p.chunk_kb);
        uint32_t accumulate = 0;
        for (; kb < kb_chunk_end; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sb = base + stage * kStageBytes;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint32_t koff = (uint32_t)k * UMMA_K * 4;
            const uint64_t dah = make_desc(sb + koff);
            const uint64_t dal = make_desc(sb + kABytes + koff);
            const uint64_t dbh = make_desc(sb + 2 * kABytes + koff);
            const uint64_t dbl = make_desc(sb + 2 * kABytes + kBBytes + koff);
            tc_mma_tf32(tmem_d, dah, dbh, kIdesc, accumulate);
            tc_mma_tf32(tmem_d, dah, dbl, kIdesc, 1u);
            if (!p.a_exact) tc_mma_tf32(tmem_d, dal, dbh, kIdesc, 1u);
            accumulate = 1u;
          }
          tc_commit(empty_bar(stage));  // frees the smem stage once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        tc_commit(tfull_bar(buf));  // accumulator chunk complete
      }
    }
  } else {
    // ================= epilogue: drain TMEM chunks into fp32 registers =================
    const int e = warp - 2;
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    const int h = e >> 2;     // column half
    float acc[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) acc[i] = 0.f;
    for (int c = 0; c < nchunks; ++c) {
      const int buf = c & 1;
      mbar_wait(tfull_bar(buf), ((uint32_t)c >> 1) & 1u);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN + h * 128);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t v[32];
        tc_ld32(taddr + j * 32, v);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[j * 32 + i] += __uint_as_float(v[i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(buf));
    }
    // ---- write the tile ----
    const int64_t m = (int64_t)tm * BM + q * 32 + lane;
    const int64_t n0 = (int64_t)tn * BN + h * 128;
    if (m < p.M) {
      if (p.mode == GEMM_SYRK) {
        float* dst = p.out + (int64_t)split * p.split_stride + m * p.ld + n0;
        // ld is a multiple of 128 and the partial buffer covers whole tiles of columns < ld
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (n0 + 4 * j < p.ld) {
            float4 v = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
            float4* d4 = reinterpret_cast<float4*>(dst) + j;
            if (p.accumulate) {
              const float4 o = *d4;
              v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
            }
            *d4 = v;
          }
        }
      } else {
        float* dst = p.out + m * p.ld + n0;
        const bool vec = (p.ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0) &&
                         ((reinterpret_cast<uintptr_t>(p.bias) & 15) == 0);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int64_t n = n0 + 4 * j;
          if (vec && n + 3 < p.N) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n));
            *(reinterpret_cast<float4*>(dst) + j) =
                make_float4(acc[4 * j] + b.x, acc[4 * j + 1] + b.y, acc[4 * j + 2] + b.z,
                            acc[4 * j + 3] + b.w);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (n + i < p.N) dst[4 * j + i] = acc[4 * j + i] + __ldg(p.bias + n + i);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512)
                 : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// operand pre-passes
// ---------------------------------------------------------------------------------------------

// SYRK operand: At[i][k] = sqrt(nw_k) (x[k0+k][i] - shift_i), split into hi / lo, transposed so
// that the sample index is contiguous.  Tile = 32 samples x 128 columns through shared memory.
__global__ void __launch_bounds__(256)
syrk_operand_kernel(const float* __restrict__ x, const float* __restrict__ nw,
                    const float* __restrict__ shift, int64_t K, int64_t D, int64_t k0, int64_t Kc,
                    int64_t Dp, float* __restrict__ at_hi, float* __restrict__ at_lo) {
  __shared__ float tile[32][129];
  const int64_t kt = (int64_t)blockIdx.x * 32;   // local sample offset in the chunk
  const int64_t it = (int64_t)blockIdx.y * 128;  // column offset
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(shift) & 15) == 0);
  for (int e = threadIdx.x; e < 32 * 32; e += 256) {
    const int r = e / 32, c4 = (e % 32) * 4;
    const int64_t k = k0 + kt + r;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (k < K && kt + r < Kc) {
      const float w = __ldg(nw + k);
      if (w != 0.f) {
        const float sw = sqrtf(w);
        const int64_t i = it + c4;
        if (vec && i + 3 < D) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(x + k * D + i));
          const float4 s = shift ? __ldg(reinterpret_cast<const float4*>(shift + i))
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
          v[0] = sw * (t.x - s.x); v[1] = sw * (t.y - s.y);
          v[2] = sw * (t.z - s.z); v[3] = sw * (t.w - s.w);
        } else {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (i + u < D) v[u] = sw * (__ldg(x + k * D + i + u) - (shift ? __ldg(shift + i + u) : 0.f));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) tile[r][c4 + u] = v[u];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < 128 * 8; e += 256) {
    const int il = e / 8, k4 = (e % 8) * 4;
    const int64_t i = it + il;
    if (i >= Dp) continue;
    float hi[4], lo[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) split_tf32(tile[k4 + u][il], hi[u], lo[u]);
    const int64_t off = i * Kc + kt + k4;
    *reinterpret_cast<float4*>(at_hi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4*>(at_lo + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// Generic row-major split: dst_hi/lo[r][c] (ld = cols_p, zero padded) from src[r][c] (ld = cols)
// or, when src == NULL, from Philox normals of the global (.., cols) base-normal matrix.
__global__ void __launch_bounds__(256)
split_rows_kernel(const float* __restrict__ src, int64_t rows, int64_t cols, int64_t cols_p,
                  uint64_t seed, uint64_t draw, uint64_t row_offset, int lower_only,
                  float* __restrict__ dst_hi, float* __restrict__ dst_lo /* NULL: src is exact */) {
  const int64_t nq_row = cols_p / 4;
  const int64_t total = rows * nq_row;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int64_t r = t / nq_row, c = (t % nq_row) * 4;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (src) {
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (c + u < cols && (!lower_only || c + u <= r)) v[u] = __ldg(src + r * cols + c + u);
    } else {
      const uint64_t e0 = (row_offset + (uint64_t)r) * (uint64_t)cols + (uint64_t)c;
      if ((cols & 3) == 0) {
        if (c < cols) philox_normal4(seed, draw, e0 >> 2, v);
      } else {
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (c + u < cols) v[u] = philox_normal1(seed, draw, e0 + u);
      }
    }
    if (dst_lo == nullptr) {
      *reinterpret_cast<float4*>(dst_hi + r * cols_p + c) = make_float4(v[0], v[1], v[2], v[3]);
      continue;
    }
    float hi[4], lo[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) split_tf32(v[u], hi[u], lo[u]);
    *reinterpret_cast<float4*>(dst_hi + r * cols_p + c) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4*>(dst_lo + r * cols_p + c) = make_float4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// cov[i][j] (+)= sum_s partial[s][min(i,j)][max(i,j)]
__global__ void syrk_reduce_kernel(const float* __restrict__ partials, int splits, int64_t Dp,
                                   int64_t split_stride, int64_t D, float* __restrict__ cov,
                                   int accumulate) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < D * D; e += stride) {
    const int64_t i = e / D, j = e % D;
    const int64_t a = i < j ? i : j, b = i < j ? j : i;
    float s = 0.f;
    for (int sp = 0; sp < splits; ++sp) s += partials[(int64_t)sp * split_stride + a * Dp + b];
    cov[e] = accumulate ? cov[e] + s : s;
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)ptr;
  }
  return fn;
}

// 2-D fp32 tensor (rows, cols) row-major with row pitch `ld` floats; box = (box_rows, BK).
static int make_map(CUtensorMap* map, const float* ptr, int64_t rows, int64_t cols, int64_t ld,
                    int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  PPI_REQUIRE(fn != nullptr, PPI_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptr, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PPI_REQUIRE(r == CUDA_SUCCESS, PPI_ERR_DRIVER, "cuTensorMapEncodeTiled failed with CUresult %d",
              (int)r);
  return 0;
}

static int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

static int launch_gemm(const CUtensorMap& ah, const CUtensorMap& al, const CUtensorMap& bh,
                       const CUtensorMap& bl, const GemmParams& p, int64_t grid, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    PPI_CUDA_TRY(cudaFuncSetAttribute(split_tf32_gemm_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)kGemmSmemBytes));
    attr_set = true;
  }
  ProfileScope prof(p.mode == GEMM_SYRK ? PPI_PROF_SYRK_GEMM : PPI_PROF_SAMPLE_GEMM, st);
  split_tf32_gemm_kernel<<<(unsigned)grid, kGemmThreads, kGemmSmemBytes, st>>>(ah, al, bh, bl, p);
  return check_launch("split_tf32_gemm_kernel");
}

constexpr int kChunkKb = 16;          // 16 k-blocks = 512 contraction steps per TMEM chunk
constexpr int64_t kSyrkKc = 65536;    // samples staged per pass
constexpr int64_t kSampleKc = 32768;  // sample rows staged per pass

struct SyrkPlan {
  int64_t Dp, Kc, tiles_m, tiles_n, ntiles;
  int splits;
  size_t op_bytes, partial_bytes;
};

static SyrkPlan syrk_plan(int64_t K, int64_t D) {
  SyrkPlan pl;
  pl.Dp = round_up(D, BM);
  pl.Kc = round_up(K < kSyrkKc ? K : kSyrkKc, (int64_t)BK * kChunkKb);
  pl.tiles_m = pl.Dp / BM;
  pl.tiles_n = (pl.Dp + BN - 1) / BN;
  pl.ntiles = 0;
  for (int64_t tm = 0; tm < pl.tiles_m; ++tm) pl.ntiles += pl.tiles_n - (tm * BM) / BN;
  int64_t s = kNumSMs / pl.ntiles;
  const int64_t max_s = pl.Kc / ((int64_t)BK * kChunkKb);
  if (s > max_s) s = max_s;
  if (s < 1) s = 1;
  pl.splits = (int)s;
  pl.op_bytes = (size_t)pl.Dp * pl.Kc * sizeof(float);
  pl.partial_bytes = (size_t)pl.splits * pl.Dp * pl.Dp * sizeof(float);
  return pl;
}

}  // namespace ppi

using namespace ppi;

extern "C" size_t ppi_wmoments_dense_workspace_bytes(int64_t K, int64_t D) {
  if (K <= 0 || D <= 0) return 0;
  const SyrkPlan pl = syrk_plan(K, D);
  return 2 * pl.op_bytes + pl.partial_bytes + 1024;
}

extern "C" int ppi_wmoments_dense(const float* x, const float* nw, int64_t K, int64_t D,
                                  const float* shift, float* cov, int accumulate, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(x && nw && cov && K > 0 && D > 0 && workspace, PPI_ERR_INVALID,
              "ppi_wmoments_dense: bad arguments");
  PPI_REQUIRE(workspace_bytes >= ppi_wmoments_dense_workspace_bytes(K, D), PPI_ERR_WORKSPACE,
              "ppi_wmoments_dense: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const SyrkPlan pl = syrk_plan(K, D);
  char* wsp = (char*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  float* at_hi = (float*)wsp;
  float* at_lo = (float*)(wsp + pl.op_bytes);
  float* partials = (float*)(wsp + 2 * pl.op_bytes);

  CUtensorMap ah, al, bh, bl;
  if (int rc = make_map(&ah, at_hi, pl.Dp, pl.Kc, pl.Kc, BM)) return rc;
  if (int rc = make_map(&al, at_lo, pl.Dp, pl.Kc, pl.Kc, BM)) return rc;
  if (int rc = make_map(&bh, at_hi, pl.Dp, pl.Kc, pl.Kc, BN)) return rc;
  if (int rc = make_map(&bl, at_lo, pl.Dp, pl.Kc, pl.Kc, BN)) return rc;

  GemmParams p;
  p.mode = GEMM_SYRK;
  p.num_kb = (int)(pl.Kc / BK);
  p.chunk_kb = kChunkKb;
  p.kb_per_split = (int)round_up((p.num_kb + pl.splits - 1) / pl.splits, kChunkKb);
  p.splits = pl.splits;
  p.tiles_m = (int)pl.tiles_m;
  p.tiles_n = (int)pl.tiles_n;
  p.M = pl.Dp;
  p.N = pl.Dp;
  p.out = partials;
  p.ld = pl.Dp;
  p.split_stride = pl.Dp * pl.Dp;
  p.bias = nullptr;
  p.a_exact = 0;

  int pass = 0;
  for (int64_t k0 = 0; k0 < K; k0 += pl.Kc, ++pass) {
    dim3 og((unsigned)(pl.Kc / 32), (unsigned)ceil_div(pl.Dp, 128));
    {
      ProfileScope prof(PPI_PROF_SYRK_OPERAND, st);
      syrk_operand_kernel<<<og, 256, 0, st>>>(x, nw, shift, K, D, k0, pl.Kc, pl.Dp, at_hi, at_lo);
      if (int rc = check_launch("syrk_operand_kernel")) return rc;
    }
    p.accumulate = pass > 0;
    if (int rc = launch_gemm(ah, al, bh, bl, p, pl.ntiles * pl.splits, st)) return rc;
  }
  const int grid = (int)(ceil_div(D * D, 256) < kNumSMs * 8 ? ceil_div(D * D, 256) : kNumSMs * 8);
  ProfileScope prof(PPI_PROF_SYRK_REDUCE, st);
  syrk_reduce_kernel<<<grid, 256, 0, st>>>(partials, pl.splits, pl.Dp, pl.Dp * pl.Dp, D, cov,
                                           accumulate);
  return check_launch("syrk_reduce_kernel");
}

// sample.cu
int ppi_sample_gaussian_small(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                              int64_t K, int D, const float* mu, const float* L, float* x,
                              cudaStream_t st);

extern "C" size_t ppi_sample_gaussian_workspace_bytes(int64_t K, int64_t D) {
  if (K <= 0 || D <= 128) return 0;
  const int64_t Dk = round_up(D, BK);
  const int64_t Kc = round_up(K < kSampleKc ? K : kSampleKc, BM);
  return 2 * (size_t)D * Dk * sizeof(float) + 2 * (size_t)Kc * Dk * sizeof(float) + 1024;
}

extern "C" int ppi_sample_gaussian(const float* z, uint64_t seed, uint64_t draw_id,
                                   uint64_t row_offset, int64_t K, int64_t D, const float* mu,
                                   const float* L, float* x, void* workspace,
                                   size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(mu && L && x && K >= 0 && D > 0, PPI_ERR_INVALID, "ppi_sample_gaussian: bad arguments");
  if (K == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (D <= 128) return ppi_sample_gaussian_small(z, seed, draw_id, row_offset, K, (int)D, mu, L, x, st);
  PPI_REQUIRE(workspace && workspace_bytes >= ppi_sample_gaussian_workspace_bytes(K, D),
              PPI_ERR_WORKSPACE, "ppi_sample_gaussian: workspace too small");
  const int64_t Dk = round_up(D, BK);
  const int64_t Kc = round_up(K < kSampleKc ? K : kSampleKc, BM);
  char* wsp = (char*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  const size_t l_bytes = (size_t)D * Dk * sizeof(float), z_bytes = (size_t)Kc * Dk * sizeof(float);
  float* l_hi = (float*)wsp;
  float* l_lo = (float*)(wsp + l_bytes);
  float* z_hi = (float*)(wsp + 2 * l_bytes);
  float* z_lo = (float*)(wsp + 2 * l_bytes + z_bytes);

  {
    const int64_t total = D * (Dk / 4);
    const int grid = (int)(ceil_div(total, 256) < kNumSMs * 8 ? ceil_div(total, 256) : kNumSMs * 8);
    split_rows_kernel<<<grid, 256, 0, st>>>(L, D, D, Dk, 0, 0, 0, 1, l_hi, l_lo);
    if (int rc = check_launch("split_rows_kernel(L)")) return rc;
  }
  CUtensorMap ah, al, bh, bl;
  if (int rc = make_map(&ah, z_hi, Kc, Dk, Dk, BM)) return rc;
  if (int rc = make_map(&al, z_lo, Kc, Dk, Dk, BM)) return rc;
  if (int rc = make_map(&bh, l_hi, D, Dk, Dk, BN)) return rc;
  if (int rc = make_map(&bl, l_lo, D, Dk, Dk, BN)) return rc;

  GemmParams p;
  p.mode = GEMM_SAMPLE;
  p.num_kb = (int)(Dk / BK);
  p.chunk_kb = 1 << 20;  // a single accumulation chunk: the contraction is only D long
  p.kb_per_split = p.num_kb;
  p.splits = 1;
  p.tiles_n = (int)ceil_div(D, BN);
  p.N = D;
  p.ld = D;
  p.split_stride = 0;
  p.bias = mu;
  p.accumulate = 0;
  p.a_exact = z ? 0 : 1;  // Philox normals sit on the TF32 grid (common.cuh box_muller)

  for (int64_t k0 = 0; k0 < K; k0 += Kc) {
    const int64_t rows = (K - k0) < Kc ? (K - k0) : Kc;
    const int64_t rows_p = round_up(rows, BM);
    const int64_t total = rows_p * (Dk / 4);
    const int grid = (int)(ceil_div(total, 256) < kNumSMs * 16 ? ceil_div(total, 256) : kNumSMs * 16);
    // rows beyond `rows` are generated too (harmless, never stored)
    {
      ProfileScope prof(PPI_PROF_SAMPLE_OPERAND, st);
      split_rows_kernel<<<grid, 256, 0, st>>>(z ? z + k0 * D : nullptr, z ? rows : rows_p, D, Dk, seed,
                                              draw_id, row_offset + (uint64_t)k0, 0, z_hi,
                                              z ? z_lo : nullptr);
      if (int rc = check_launch("split_rows_kernel(z)")) return rc;
    }
    p.tiles_m = (int)(rows_p / BM);
    p.M = rows;
    p.out = x + k0 * D;
    if (int rc = launch_gemm(ah, al, bh, bl, p, (int64_t)p.tiles_m * p.tiles_n, st)) return rc;
  }
  return 0;
}
