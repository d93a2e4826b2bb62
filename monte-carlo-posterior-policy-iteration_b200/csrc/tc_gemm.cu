// tcgen05 / TMA kernels of the PPI update path (sm_100a).
//
//   ppi_wmoments_dense   sum_k nw_k (x_k - s)(x_k - s)^T          ppi/policies.py:89-92
//   ppi_sample_gaussian  x = mu + z L^T  (D > 128)                ppi/policies.py:247-249,
//                                                                  ppi/samplers.py:38-39
//
// Both are tensor-core contractions  C[m,n] = sum_k A[m,k] B[n,k]  with both operands stored
// K-major (contraction index contiguous), the canonical tcgen05 shared-memory layout (128-byte
// swizzle, 8-row atoms), staged by operand pre-passes and fed by TMA.
//
// Precision.  tcgen05 has no fp32-input MMA, and its fp32 accumulation TRUNCATES: measured here,
// every MMA instruction that adds into a TMEM accumulator costs ~3e-8 relative (biased), so a
// chain of n instructions drifts by ~3e-8 n.  Two consequences shape both kernels:
//   (1) fp32 operand values are split by a pre-pass into exact tensor-core operands and the
//       product is rebuilt from partial products (scheme chosen with ppi_set_option):
//         PPI_GEMM_FP16X2  v 2^e = h1 + h2 + O(2^-22 |v 2^e|), h1 = fp16(v 2^e), h2 = fp16(rest); one
//                          power-of-two scale 2^e per operand puts its largest entry in
//                          [2^13, 2^14), so entries down to 2^-28 of the largest keep 22 bits.
//                          h1 g1 + h1 g2 + h2 g1 as three kind::f16 MMAs: error <= 3 2^-22 |a||b|,
//                          4 bytes per operand value, 6 MMA instructions per 32-deep k-block.
//         PPI_GEMM_TF32X3  hi = tf32(v), lo = tf32(v - hi); hi hi + hi lo + lo hi as three
//                          kind::tf32 MMAs: error ~2^-21 |a||b| at any magnitude, 8 bytes per
//                          value, 12 MMA instructions per k-block.  Kept as the conservative path.
//       Philox normals are exact single fp16 / tf32 operands (common.cuh box_muller).
//   (2) the TMEM accumulator is drained into fp32 registers (round-to-nearest adds) every
//       `chunk` k-blocks -- two 256-column TMEM buffers ping-pong, so draining overlaps the MMAs
//       of the next chunk -- which bounds every in-tensor-core chain to a few dozen instructions.
// Both kernels are fed faster by the tensor pipe than L2 can deliver operands (ncu: ~54 B/clk/SM
// of TMA ingest), which is why the 4-byte operand form matters more than the MMA count.
//
// CTA = 10 warps: warp 0 TMA producer (one lane), warp 1 MMA issuer (one lane) + TMEM
// allocation, warps 2..9 drain TMEM (each warp owns its hardware lane quarter and one 128-column
// half).  One 128 x 256 accumulator per TMEM buffer.
//
// Covariance: one upper-triangular 128 x 256 output tile x one K-split per CTA, 144 CTAs.
// Sampling: transposed product  X^T[j, s] = sum_k L[j,k] z[s,k]  so that TMEM lanes are output
// columns j and a warp's 32 lanes store 128 contiguous bytes of one sample row; persistent CTAs
// walk (sample block, j-tile pair) work units -- pairs (i, T-1-i) have equal triangular cost --
// and the epilogue of one tile overlaps the MMAs of the next through the two TMEM buffers.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace ppi {

constexpr int BM = 128;        // UMMA M (TMEM lanes)
constexpr int BN = 256;        // UMMA N (TMEM columns per accumulator)
constexpr int BK = 32;         // contraction steps per k-block (one 128-byte row of fp32 or of an fp16 pair)
constexpr int kGemmThreads = 320;
constexpr uint32_t kABytes = BM * 128;      // 16 KB: 128 rows x 128 B
constexpr uint32_t kBBytes = BN * 128;      // 32 KB
constexpr int kMaxStages = 4;
constexpr uint32_t kPipeBytes = 192 * 1024; // every variant uses stages * stage_bytes = 192 KB
constexpr uint32_t kGemmSmemBytes = kPipeBytes + 1024 /*align*/ + 256 /*barriers*/;

// K-major base-normal store (ppi_sample_gaussian_keep, Philox): blocks of kZBlock samples
constexpr int64_t kZBlock = 2048;
__host__ __device__ inline int64_t zt_index(int64_t k, int64_t i, int64_t rows) {
  return ((k / kZBlock) * rows + i) * kZBlock + (k % kZBlock);
}

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

// kind::f16 with fp16 inputs, fp32 accumulate, A and B K-major.
constexpr uint32_t kIdescF16 = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// ---- split products of one pipeline stage into the accumulator at tmem_d -------------------------
// TF32X3 stage: A_hi | A_lo | B_hi | B_lo (fp32 rows of 32), 32 contraction steps.
__device__ __forceinline__ void issue_tf32x3(uint32_t tmem_d, uint32_t sb, uint32_t& accumulate,
                                             bool use_alo, bool use_blo) {
  const uint32_t ah = sb, al = sb + kABytes, bh = sb + 2 * kABytes, bl = sb + 2 * kABytes + kBBytes;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t koff = (uint32_t)k * 32u;  // 8 tf32 = 32 bytes inside the 128-byte swizzle row
    const uint64_t dah = make_desc(ah + koff), dbh = make_desc(bh + koff);
    tc_mma_tf32(tmem_d, dah, dbh, kIdesc, accumulate);
    accumulate = 1u;
    if (use_blo) tc_mma_tf32(tmem_d, dah, make_desc(bl + koff), kIdesc, 1u);
    if (use_alo) tc_mma_tf32(tmem_d, make_desc(al + koff), dbh, kIdesc, 1u);
  }
}
// FP16X2 pair x pair: rows are [h1 x 32 | h2 x 32] fp16, 32 contraction steps, 2 k-steps of 16.
__device__ __forceinline__ void issue_f16_pair_pair(uint32_t tmem_d, uint32_t a, uint32_t b,
                                                    uint32_t& accumulate) {
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const uint32_t off = (uint32_t)k * 32u;  // 16 fp16 = 32 bytes
    const uint64_t a1 = make_desc(a + off), b1 = make_desc(b + off);
    tc_mma_f16(tmem_d, a1, b1, kIdescF16, accumulate);
    accumulate = 1u;
    tc_mma_f16(tmem_d, a1, make_desc(b + 64u + off), kIdescF16, 1u);
    tc_mma_f16(tmem_d, make_desc(a + 64u + off), b1, kIdescF16, 1u);
  }
}
// FP16X2 pair x single: A rows [h1 x 32 | h2 x 32], B rows hold 64 exact fp16 (two k-blocks);
// `half` selects which 32 of them this A box multiplies.
__device__ __forceinline__ void issue_f16_pair_single(uint32_t tmem_d, uint32_t a, uint32_t b, int half,
                                                      uint32_t& accumulate) {
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const uint32_t off = (uint32_t)k * 32u;
    const uint64_t bz = make_desc(b + (uint32_t)half * 64u + off);
    tc_mma_f16(tmem_d, make_desc(a + off), bz, kIdescF16, accumulate);
    accumulate = 1u;
    tc_mma_f16(tmem_d, make_desc(a + 64u + off), bz, kIdescF16, 1u);
  }
}

// MN-major operand, 128-byte swizzle (cute::UMMA make_umma_desc<Major::MN>): rows of 128 B = 64 consecutive
// MN-elements, one row per contraction index; 8 rows = one 1024-byte swizzle atom.  LBO = bytes between
// 64-element chunks along MN, SBO = bytes between groups of 8 contraction rows.  This is how a TMA box
// [64 columns i][64 samples] of the K-major base-normal store lands in shared memory when the SAMPLING GEMM
// (contraction over i, N = samples) reads it: one such box per 64 samples, 8 KB apart.
constexpr uint32_t kZBoxBytes = 64 * 128;   // [64 contraction rows][64 samples] fp16
__device__ __forceinline__ uint64_t make_desc_mn(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)(kZBoxBytes >> 4) << 16;   // LBO: next 64 samples = next box
  d |= (uint64_t)(1024 >> 4) << 32;         // SBO: next 8 contraction rows
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16, fp32 accumulate, A K-major, B MN-major (b_major bit 16)
constexpr uint32_t kIdescF16Bmn = kIdescF16 | (1u << 16);

// FP16X2 pair x single with B = base normals straight from the K-major store (MN-major for this product):
// A rows [h1 x 32 | h2 x 32] of the 32-deep k-block `half` of the stage; B = four boxes of 64 samples,
// contraction rows 32 half .. 32 half + 31 of the stage (16 per MMA: two swizzle atoms, 2 KB).
__device__ __forceinline__ void issue_f16_pair_store(uint32_t tmem_d, uint32_t a, uint32_t b, int half,
                                                     uint32_t& accumulate) {
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const uint32_t off = (uint32_t)k * 32u;
    const uint64_t bz = make_desc_mn(b + (uint32_t)(half * 2 + k) * 2048u);
    tc_mma_f16(tmem_d, make_desc(a + off), bz, kIdescF16Bmn, accumulate);
    accumulate = 1u;
    tc_mma_f16(tmem_d, make_desc(a + 64u + off), bz, kIdescF16Bmn, 1u);
  }
}

template <int N>
__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }

// ---- shared CTA prologue: barriers + TMEM allocation ----------------------------------------------
struct GemmSmem {
  uint32_t base;      // 1024-aligned shared address of stage 0
  uint32_t bar_base;  // barriers: full[kMaxStages], empty[kMaxStages], tfull[2], tempty[2]
  __device__ __forceinline__ uint32_t full_bar(int s) const { return bar_base + 8u * s; }
  __device__ __forceinline__ uint32_t empty_bar(int s) const { return bar_base + 8u * (kMaxStages + s); }
  __device__ __forceinline__ uint32_t tfull_bar(int b) const { return bar_base + 8u * (2 * kMaxStages + b); }
  __device__ __forceinline__ uint32_t tempty_bar(int b) const { return bar_base + 8u * (2 * kMaxStages + 2 + b); }
};

__device__ __forceinline__ uint32_t gemm_prologue(uint8_t* smem_raw, GemmSmem& gs, int epi_warps = 8) {
  const uint32_t raw = smem_u32(smem_raw);
  gs.base = (raw + 1023u) & ~1023u;
  gs.bar_base = gs.base + kPipeBytes;
  uint8_t* smem = smem_raw + (gs.base - raw);
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem + kPipeBytes + 8 * (2 * kMaxStages + 4));
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kMaxStages; ++s) { mbar_init(gs.full_bar(s), 1); mbar_init(gs.empty_bar(s), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(gs.tfull_bar(b), 1); mbar_init(gs.tempty_bar(b), epi_warps); }
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
  return *tmem_ptr_smem;
}

__device__ __forceinline__ void gemm_teardown(uint32_t tmem_base) {
  tc_fence_before();
  __syncthreads();
  if ((threadIdx.x >> 5) == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512)
                 : "memory");
  }
}

// acc[128] += the 128 columns [col0, col0+128) of this warp's TMEM lane quarter
__device__ __forceinline__ void drain_add(uint32_t taddr, float (&acc)[128]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t v[32];
    tc_ld32(taddr + j * 32, v);
    tc_wait_ld();
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[j * 32 + i] += __uint_as_float(v[i]);
  }
}

// acc[64] += the 64 columns [col0, col0+64) of this warp's TMEM lane quarter
__device__ __forceinline__ void drain_add64(uint32_t taddr, float (&acc)[64]) {
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    uint32_t v[32];
    tc_ld32(taddr + j * 32, v);
    tc_wait_ld();
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[j * 32 + i] += __uint_as_float(v[i]);
  }
}

// Device-side scalars of one GEMM call (filled by tiny kernels, never read by the host).
struct GemmScalars {
  unsigned int amax_a_bits;  // max |.| of operand A source (non-negative float bits, atomicMax)
  unsigned int amax_b_bits;
  float scale_a, scale_b;    // power-of-two operand scales (1 for TF32X3)
  float inv_scale;           // 1 / (scale_a * scale_b)
  float pad[3];
};

// ---------------------------------------------------------------------------------------------
// Covariance kernel: partial[split][tile] (+)= A_tile A_tile'^T over this CTA's k-range
// ---------------------------------------------------------------------------------------------
struct SyrkParams {
  int num_kb;        // k-blocks in the staged operand
  int chunk_kb;      // k-blocks per TMEM accumulation chunk
  int kb_per_split;  // k-blocks handled by one CTA
  int splits;        // CTAs per output tile
  int tiles_n;
  float* out;        // partials [splits][ld][ld]
  int64_t ld;
  int64_t split_stride;
  int accumulate;    // partial += acc
};

// MODE (one "k-unit" = one pipeline stage):
//   SYRK_TF32X3       unit = 32 steps: A_hi | A_lo | B_hi | B_lo (fp32 rows),               2 stages
//   SYRK_PAIR         unit = 32 steps: A pair (16 KB) | B pair (32 KB),                      4 stages
//   SYRK_PAIR_SINGLE  unit = 64 steps: A pair x2 (32 KB) | B exact single fp16 (32 KB),      3 stages
//                     (whitened moments: A = nw z as a pair, B = z, 2 partial products)
enum { SYRK_TF32X3 = 0, SYRK_PAIR = 1, SYRK_PAIR_SINGLE = 2 };

template <int MODE>
__global__ void __launch_bounds__(kGemmThreads, 1)
syrk_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_a2,
                 const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_b2,
                 const SyrkParams p) {
  constexpr bool FP16 = MODE != SYRK_TF32X3;
  constexpr bool kSingle = MODE == SYRK_PAIR_SINGLE;
  constexpr int kStages = kSingle ? 3 : (FP16 ? 4 : 2);
  constexpr uint32_t kStageBytes = kSingle ? (2 * kABytes + kBBytes)
                                           : (FP16 ? (kABytes + kBBytes) : 2 * (kABytes + kBBytes));
  static_assert(kStages * kStageBytes == kPipeBytes, "pipeline must fill the smem budget");
  constexpr int kx = FP16 ? 2 * BK : BK;  // elements per 32-step k-block in a pair / fp32 operand array

  extern __shared__ uint8_t smem_raw[];
  GemmSmem gs;
  const uint32_t tmem_base = gemm_prologue(smem_raw, gs);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- tile / k-range of this CTA: upper-triangular tiles, tn >= tn_min(tm) ----
  int t = blockIdx.x / p.splits;
  const int split = blockIdx.x % p.splits;
  int tm = 0, tn = 0;
  for (;; ++tm) {
    const int tn_min = (tm * BM) / BN;
    const int cnt = p.tiles_n - tn_min;
    if (t < cnt) { tn = tn_min + t; break; }
    t -= cnt;
  }
  const int kb_begin = split * p.kb_per_split;
  const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
  const int nkb = max(0, kb_end - kb_begin);
  const int nchunks = (nkb + p.chunk_kb - 1) / p.chunk_kb;

  if (warp == 0) {
    if (lane == 0 && nkb > 0) {
      // ================= TMA producer =================
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
      if (!FP16) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a2) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b2) : "memory");
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(gs.empty_bar(stage), phase ^ 1u);
        const uint32_t sb = gs.base + stage * kStageBytes;
        mbar_expect_tx(gs.full_bar(stage), kStageBytes);
        if (kSingle) {
          tma_load_2d(sb, &map_a, gs.full_bar(stage), (2 * kb) * kx, tm * BM);
          tma_load_2d(sb + kABytes, &map_a, gs.full_bar(stage), (2 * kb + 1) * kx, tm * BM);
          tma_load_2d(sb + 2 * kABytes, &map_b, gs.full_bar(stage), kb * 64, tn * BN);  // 64 exact fp16 per row
        } else if (FP16) {
          tma_load_2d(sb, &map_a, gs.full_bar(stage), kb * kx, tm * BM);
          tma_load_2d(sb + kABytes, &map_b, gs.full_bar(stage), kb * kx, tn * BN);
        } else {
          tma_load_2d(sb, &map_a, gs.full_bar(stage), kb * kx, tm * BM);
          tma_load_2d(sb + kABytes, &map_a2, gs.full_bar(stage), kb * kx, tm * BM);
          tma_load_2d(sb + 2 * kABytes, &map_b, gs.full_bar(stage), kb * kx, tn * BN);
          tma_load_2d(sb + 2 * kABytes + kBBytes, &map_b2, gs.full_bar(stage), kb * kx, tn * BN);
        }
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
        mbar_wait(gs.tempty_bar(buf), (((uint32_t)c >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)buf * BN;
        const int kb_chunk_end = min(nkb, kb + p.chunk_kb);
        uint32_t accumulate = 0;
        for (; kb < kb_chunk_end; ++kb) {
          mbar_wait(gs.full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sb = gs.base + stage * kStageBytes;
          if (kSingle) {
            issue_f16_pair_single(tmem_d, sb, sb + 2 * kABytes, 0, accumulate);
            issue_f16_pair_single(tmem_d, sb + kABytes, sb + 2 * kABytes, 1, accumulate);
          } else if (FP16) {
            issue_f16_pair_pair(tmem_d, sb, sb + kABytes, accumulate);
          } else {
            issue_tf32x3(tmem_d, sb, accumulate, true, true);
          }
          tc_commit(gs.empty_bar(stage));  // frees the smem stage once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        tc_commit(gs.tfull_bar(buf));  // accumulator chunk complete
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
      mbar_wait(gs.tfull_bar(buf), ((uint32_t)c >> 1) & 1u);
      tc_fence_after();
      drain_add(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN + h * 128), acc);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(gs.tempty_bar(buf));
    }
    // ---- write the tile (ld is a multiple of 128; the partial buffer covers whole tiles) ----
    const int64_t m = (int64_t)tm * BM + q * 32 + lane;
    const int64_t n0 = (int64_t)tn * BN + h * 128;
    if (m < p.ld) {
      float* dst = p.out + (int64_t)split * p.split_stride + m * p.ld + n0;
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
    }
  }
  gemm_teardown(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// Sampling kernel (transposed, persistent):  x[s, j] = mu[j] + sum_{k<=j} L[j,k] z[s,k]
//   A = L rows (UMMA M = 128 output columns j -> TMEM lanes), B = z rows (UMMA N = 256 samples
//   -> TMEM columns).  Work unit = (sample block nb, pair p) -> j-tiles T-1-p then p.
//   FP16, !ZPAIR  Philox z = one exact fp16 per value: stage = 64 steps: L pair x2 | z single, 3 stages
//   FP16,  ZPAIR  injected z = fp16 pair:              stage = 32 steps: L pair | z pair,      4 stages
//   !FP16         TF32X3:                               stage = 32 steps: L_hi|L_lo|z_hi|z_lo,  2 stages
// ---------------------------------------------------------------------------------------------
struct SampleParams {
  int tiles_j;          // ceil(D / 128)
  int units_per_block;  // ceil(tiles_j / 2)
  int num_st;           // pipeline stages covering the padded contraction length
  int chunk_st;         // stages per TMEM accumulation chunk
  int64_t n_units;      // sample blocks * units_per_block
  int64_t rows;         // valid samples in this pass
  int64_t D;
  float* x;             // [rows][D]
  const float* mu;
  const GemmScalars* scalars;  // inv_scale (FP16) or NULL
  int z_exact;          // TF32X3 only: z is exactly tf32 (Philox), its low-order operand is skipped
  // fused separable cost (optimization/functions.py:65-66,78-83): per-(sample, j-tile) partial sums of the
  // cost terms of the 128 columns of the tile, written to cost_part[(row0 + s) * tiles_j + tj]; x may then be
  // NULL (nobody reads the samples: the whitened update contracts the stored base normals)
  float* cost_part;
  int cost_fn;          // PPI_FN_STYBLINSKI / PPI_FN_RASTRIGIN
  int64_t row0;         // first sample row of this pass inside cost_part
  // Philox normals (FP16, !ZPAIR): map_z is the blocked K-major store zT[k / 2048][i][k % 2048]
  int64_t store_k0;     // store sample index of this pass' first sample (multiple of 256)
  int store_rows;       // rows per 2048-sample block of the store (round_up(D, 128))
};

// term of a separable test function at one coordinate (cost.cu has the same expressions)
__device__ __forceinline__ float fused_cost_term(int fn, float x) {
  if (fn == PPI_FN_STYBLINSKI) {
    const float x2 = x * x;
    return fmaf(x2, x2, fmaf(-16.f, x2, 5.f * x));
  }
  const float fr = x - rintf(x);   // Rastrigin: exact range reduction, one MUFU.COS
  return fmaf(x, x, -10.f * __cosf(6.283185307179586f * fr));
}

// Transpose-reduce across the warp: every lane holds t[0..63] (one value per sample, its own column); on
// return lane l holds in t[0..1] the sums over the 32 lanes for samples 2 l, 2 l + 1.  62 shuffles.
__device__ __forceinline__ void warp_transpose_reduce64(float (&t)[64]) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int w = 32, off = 16; w >= 2; w >>= 1, off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      if (i < w) {
        const float send = up ? t[i] : t[i + w];
        const float keep = up ? t[i + w] : t[i];
        t[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
      }
    }
  }
}

// CTA = 5 warpgroups: WG0 = warp 0 TMA producer, warp 1 MMA issuer (+ two idle warps), WG1-4 = sixteen warps
// that drain TMEM: warp w owns its hardware lane quarter (w & 3) and one 64-column (sample) group -- sixteen
// short epilogue tails instead of eight long ones, so that the per-tile tail (bias, optional store, optional
// cost terms + their reduction over the tile's columns) stays inside the two accumulator chunks the MMA issuer
// can run ahead.  The CTA's registers are re-divided with setmaxnreg (WG0 40, drain warps 104 per thread).
constexpr int kSampleThreads = 640;
// setmaxnreg re-divides the CTA's OWN allocation (registers per thread at launch x threads; ptxas gives a
// 640-thread CTA at most 96), not the SM's 64 K: a request beyond it blocks in setmaxnreg.inc for ever.
constexpr int kSampleRegsLaunch = 96, kSampleRegsIssue = 40, kSampleRegsDrain = 104;
static_assert(4 * 32 * kSampleRegsIssue + 16 * 32 * kSampleRegsDrain <= kSampleRegsLaunch * kSampleThreads,
              "setmaxnreg budget of sample_gemm_kernel exceeds the CTA's register allocation");
template <bool FP16, bool ZPAIR>
__global__ void __launch_bounds__(kSampleThreads, 1)
sample_gemm_kernel(const __grid_constant__ CUtensorMap map_l, const __grid_constant__ CUtensorMap map_l2,
                   const __grid_constant__ CUtensorMap map_z, const __grid_constant__ CUtensorMap map_z2,
                   const SampleParams p) {
  constexpr bool kSingle = FP16 && !ZPAIR;
  constexpr int KS = kSingle ? 64 : 32;  // contraction steps per stage
  constexpr int kStages = kSingle ? 3 : (FP16 ? 4 : 2);
  constexpr uint32_t kStageBytes = kSingle ? (2 * kABytes + kBBytes)
                                           : (FP16 ? (kABytes + kBBytes) : 2 * (kABytes + kBBytes));
  static_assert(kStages * kStageBytes == kPipeBytes, "pipeline must fill the smem budget");

  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(16) float cost_red[2][4][4][64];   // [tile parity][sample group][lane quarter][sample]
  GemmSmem gs;
  const uint32_t tmem_base = gemm_prologue(smem_raw, gs, 16);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool load_zlo = !FP16 && !p.z_exact;

  // Every role walks the same tile sequence.
  auto tile_j = [&](int pr, int tt) { return tt == 0 ? p.tiles_j - 1 - pr : pr; };
  auto tiles_in_unit = [&](int pr) { return (p.tiles_j - 1 - pr != pr) ? 2 : 1; };
  auto stages_of = [&](int tj) { return min(p.num_st, ((tj + 1) * BM + KS - 1) / KS); };  // L lower triangular

  if (warp < 4) {
  reg_dec<kSampleRegsIssue>();
  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_l) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_z) : "memory");
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t u = blockIdx.x; u < p.n_units; u += gridDim.x) {
        const int nb = (int)(u / p.units_per_block), pr = (int)(u % p.units_per_block);
        for (int tt = 0; tt < tiles_in_unit(pr); ++tt) {
          const int tj = tile_j(pr, tt), nst = stages_of(tj);
          for (int s = 0; s < nst; ++s) {
            mbar_wait(gs.empty_bar(stage), phase ^ 1u);
            const uint32_t sb = gs.base + stage * kStageBytes;
            const uint32_t fb = gs.full_bar(stage);
            if (kSingle) {
              mbar_expect_tx(fb, kStageBytes);
              tma_load_2d(sb, &map_l, fb, (2 * s) * 64, tj * BM);            // L pair, k-block 2s
              tma_load_2d(sb + kABytes, &map_l, fb, (2 * s + 1) * 64, tj * BM);  // k-block 2s+1
              // base normals of columns 64 s .. 64 s + 63 for the block's 256 samples: four boxes
              // [64 store rows][64 samples] of the K-major store (MN-major operand of this product)
              const int64_t smp = p.store_k0 + (int64_t)nb * BN;
              const int c0 = (int)(smp % kZBlock), rb = (int)(smp / kZBlock) * p.store_rows + s * 64;
#pragma unroll
              for (int b = 0; b < 4; ++b)
                tma_load_2d(sb + 2 * kABytes + b * kZBoxBytes, &map_z, fb, c0 + 64 * b, rb);
            } else if (FP16) {
              mbar_expect_tx(fb, kStageBytes);
              tma_load_2d(sb, &map_l, fb, s * 64, tj * BM);
              tma_load_2d(sb + kABytes, &map_z, fb, s * 64, nb * BN);
            } else {
              mbar_expect_tx(fb, load_zlo ? kStageBytes : kStageBytes - kBBytes);
              tma_load_2d(sb, &map_l, fb, s * BK, tj * BM);
              tma_load_2d(sb + kABytes, &map_l2, fb, s * BK, tj * BM);
              tma_load_2d(sb + 2 * kABytes, &map_z, fb, s * BK, nb * BN);
              if (load_zlo) tma_load_2d(sb + 2 * kABytes + kBBytes, &map_z2, fb, s * BK, nb * BN);
            }
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ================= MMA issuer =================
      int stage = 0;
      uint32_t phase = 0, ccount = 0;  // ccount: chunks issued so far (selects the TMEM buffer)
      for (int64_t u = blockIdx.x; u < p.n_units; u += gridDim.x) {
        const int pr = (int)(u % p.units_per_block);
        for (int tt = 0; tt < tiles_in_unit(pr); ++tt) {
          const int nst = stages_of(tile_j(pr, tt));
          for (int s0 = 0; s0 < nst; s0 += p.chunk_st, ++ccount) {
            const uint32_t buf = ccount & 1u;
            mbar_wait(gs.tempty_bar(buf), ((ccount >> 1) & 1u) ^ 1u);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + buf * BN;
            uint32_t accumulate = 0;
            const int s1 = min(nst, s0 + p.chunk_st);
            for (int s = s0; s < s1; ++s) {
              mbar_wait(gs.full_bar(stage), phase);
              tc_fence_after();
              const uint32_t sb = gs.base + stage * kStageBytes;
              if (kSingle) {
                issue_f16_pair_store(tmem_d, sb, sb + 2 * kABytes, 0, accumulate);
                issue_f16_pair_store(tmem_d, sb + kABytes, sb + 2 * kABytes, 1, accumulate);
              } else if (FP16) {
                issue_f16_pair_pair(tmem_d, sb, sb + kABytes, accumulate);
              } else {
                issue_tf32x3(tmem_d, sb, accumulate, true, !p.z_exact);
              }
              tc_commit(gs.empty_bar(stage));
              if (++stage == kStages) { stage = 0; phase ^= 1u; }
            }
            tc_commit(gs.tfull_bar(buf));
          }
        }
      }
    }
  }
  } else {
    // setmaxnreg moves registers inside the CTA's own allocation (96 x 640 at launch): 40 x 128 + 104 x 512 fits
    reg_inc<kSampleRegsDrain>();
    // ================= epilogue: TMEM -> registers -> global, 128 B per warp store =================
    const int e = warp - 4;
    const int q = warp & 3;   // TMEM lane quarter
    const int h = e >> 2;     // sample (column) group of 64
    const float inv_scale = p.scalars ? p.scalars->inv_scale : 1.0f;
    uint32_t ccount = 0, tile_count = 0;
    for (int64_t u = blockIdx.x; u < p.n_units; u += gridDim.x) {
      const int nb = (int)(u / p.units_per_block), pr = (int)(u % p.units_per_block);
      for (int tt = 0; tt < tiles_in_unit(pr); ++tt) {
        const int tj = tile_j(pr, tt), nst = stages_of(tj);
        const int64_t j = (int64_t)tj * BM + q * 32 + lane;
        const bool jok = j < p.D;
        const float bias = jok ? __ldg(p.mu + j) : 0.f;
        float acc[64];
#pragma unroll
        for (int i = 0; i < 64; ++i) acc[i] = 0.f;
        for (int s0 = 0; s0 < nst; s0 += p.chunk_st, ++ccount) {
          const uint32_t buf = ccount & 1u;
          mbar_wait(gs.tfull_bar(buf), (ccount >> 1) & 1u);
          tc_fence_after();
          drain_add64(tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN + (uint32_t)(h * 64), acc);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(gs.tempty_bar(buf));
        }
        const int64_t s_first = (int64_t)nb * BN + h * 64;
        const int64_t left = p.rows - s_first;
        if (left > 0) {                      // uniform over the four quarter warps of this sample group
#pragma unroll
          for (int c = 0; c < 64; ++c) acc[c] = fmaf(acc[c], inv_scale, bias);
          if (jok && p.x != nullptr) {
            float* dst = p.x + s_first * p.D + j;
            if (left >= 64) {
#pragma unroll
              for (int c = 0; c < 64; ++c) dst[(int64_t)c * p.D] = acc[c];
            } else {
#pragma unroll
              for (int c = 0; c < 64; ++c)
                if (c < left) dst[(int64_t)c * p.D] = acc[c];
            }
          }
          if (p.cost_part != nullptr) {
            // cost terms of this thread's column for its 64 samples -> sums over the tile's 128 columns:
            // across the 32 lanes by the transpose-reduce, across the four quarter warps through shared memory
            // (double-buffered by tile parity: one named barrier per tile), in a FIXED order (deterministic)
#pragma unroll
            for (int c = 0; c < 64; ++c) acc[c] = jok ? fused_cost_term(p.cost_fn, acc[c]) : 0.f;
            warp_transpose_reduce64(acc);
            *reinterpret_cast<float2*>(&cost_red[tile_count & 1][h][q][2 * lane]) = make_float2(acc[0], acc[1]);
            asm volatile("bar.sync %0, %1;" ::"r"(1 + h), "r"(128) : "memory");
            if (q == 0) {
              float2 r = *reinterpret_cast<const float2*>(&cost_red[tile_count & 1][h][0][2 * lane]);
#pragma unroll
              for (int qq = 1; qq < 4; ++qq) {
                const float2 o = *reinterpret_cast<const float2*>(&cost_red[tile_count & 1][h][qq][2 * lane]);
                r.x += o.x; r.y += o.y;
              }
              if (2 * lane < left) p.cost_part[(p.row0 + s_first + 2 * lane) * p.tiles_j + tj] = r.x;
              if (2 * lane + 1 < left) p.cost_part[(p.row0 + s_first + 2 * lane + 1) * p.tiles_j + tj] = r.y;
            }
          }
        }
        ++tile_count;
      }
    }
  }
  gemm_teardown(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// 2-CTA helpers (cta_group::2): a cluster of two CTAs on one TPC issues one UMMA of M = 256; every
// CTA keeps its own 128 A rows and HALF of the B rows in its shared memory, so the shared-memory
// traffic per MMA (the resource that bounds the 1-CTA kernels: 12 KB of operand reads per 128-cycle
// MMA + the TMA fill, against 128 B/clk) drops by a third.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `addr` (a shared::cta address of this CTA) in CTA `rank`
__device__ __forceinline__ uint32_t mapa_rank(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
// TMA load whose completion bytes are counted on the LEADER CTA's mbarrier (peer bit cleared)
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                                int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_commit_2sm(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_f16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// kind::f16, fp32 accumulate, K-major operands, M = 256 (two CTAs x 128 rows), N = 256
constexpr uint32_t kIdescF16Pair = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);

// pair x single over one 64-step stage: A = two pair tiles (32 steps each), B = this CTA's half of the
// 256 B rows (128 rows x 64 exact fp16)
__device__ __forceinline__ void issue_f16_pair_single_2sm(uint32_t tmem_d, uint32_t a, uint32_t b, int half,
                                                          uint32_t& accumulate) {
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const uint32_t off = (uint32_t)k * 32u;
    const uint64_t bz = make_desc(b + (uint32_t)half * 64u + off);
    tc_mma_f16_2sm(tmem_d, make_desc(a + off), bz, kIdescF16Pair, accumulate);
    accumulate = 1u;
    tc_mma_f16_2sm(tmem_d, make_desc(a + 64u + off), bz, kIdescF16Pair, 1u);
  }
}

// ---------------------------------------------------------------------------------------------
// operand pre-passes
// ---------------------------------------------------------------------------------------------
enum { OP_TF32X3 = 0, OP_F16_PAIR = 1, OP_F16_SINGLE = 2 };

__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
  const __half2 t = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&t);
}

// Store the split of 4 consecutive contraction elements (k .. k+3, k % 4 == 0) of operand row r;
// v is already multiplied by the operand's power-of-two scale.  kp = padded contraction length.
//   OP_TF32X3     op_a [rows][kp] fp32 hi = tf32(v);  op_b [rows][kp] fp32 lo = tf32(v - hi) (may be NULL)
//   OP_F16_PAIR   op_a [rows][kp/32][ h1 x 32 | h2 x 32 ] fp16, h1 = fp16(v), h2 = fp16(v - h1)
//   OP_F16_SINGLE op_a [rows][kp] fp16 (v is exactly representable: Philox normals)
template <int MODE>
__device__ __forceinline__ void store_split4(const float v[4], int64_t r, int64_t k, int64_t kp,
                                             void* __restrict__ op_a, void* __restrict__ op_b) {
  if (MODE == OP_TF32X3) {
    float hi[4], lo[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) split_tf32(v[u], hi[u], lo[u]);
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(op_a) + r * kp + k) =
        make_float4(hi[0], hi[1], hi[2], hi[3]);
    if (op_b)
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(op_b) + r * kp + k) =
          make_float4(lo[0], lo[1], lo[2], lo[3]);
  } else if (MODE == OP_F16_PAIR) {
    float r2[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) r2[u] = v[u] - __half2float(__float2half_rn(v[u]));  // exact in fp32
    uint16_t* base = reinterpret_cast<uint16_t*>(op_a) + r * (2 * kp) + (k >> 5) * 64 + (k & 31);
    *reinterpret_cast<uint2*>(base) = make_uint2(pack_half2(v[0], v[1]), pack_half2(v[2], v[3]));
    *reinterpret_cast<uint2*>(base + 32) = make_uint2(pack_half2(r2[0], r2[1]), pack_half2(r2[2], r2[3]));
  } else {
    uint16_t* base = reinterpret_cast<uint16_t*>(op_a) + r * kp + k;
    *reinterpret_cast<uint2*>(base) = make_uint2(pack_half2(v[0], v[1]), pack_half2(v[2], v[3]));
  }
}

// ---- power-of-two operand scales (FP16X2) -----------------------------------------------------------
__device__ __forceinline__ void atomic_absmax(unsigned int* out, float m) {
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out, __float_as_uint(m));  // non-negative floats order as uints
}

__global__ void __launch_bounds__(256) absmax_kernel(const float* __restrict__ a, int64_t n,
                                                     unsigned int* __restrict__ out) {
  float m = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    m = fmaxf(m, fabsf(__ldg(a + i)));
  atomic_absmax(out, m);
}

// max |x[k][i] - shift[i]| over rows with non-zero weight
__global__ void __launch_bounds__(256)
absmax_dev_kernel(const float* __restrict__ x, const float* __restrict__ nw,
                  const float* __restrict__ shift, int64_t K, int64_t D, unsigned int* __restrict__ out) {
  float m = 0.f;
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(shift) & 15) == 0);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {
    const int64_t nq = D / 4;
    for (int64_t t = tid; t < K * nq; t += stride) {
      const int64_t k = t / nq, i = (t % nq) * 4;
      if (__ldg(nw + k) == 0.f) continue;
      const float4 v = __ldg(reinterpret_cast<const float4*>(x + k * D + i));
      const float4 s = shift ? __ldg(reinterpret_cast<const float4*>(shift + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
      m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x - s.x), fabsf(v.y - s.y))), fmaxf(fabsf(v.z - s.z), fabsf(v.w - s.w)));
    }
  } else {
    for (int64_t t = tid; t < K * D; t += stride) {
      const int64_t k = t / D, i = t % D;
      if (__ldg(nw + k) == 0.f) continue;
      m = fmaxf(m, fabsf(__ldg(x + t) - (shift ? __ldg(shift + i) : 0.f)));
    }
  }
  atomic_absmax(out, m);
}

// 2^e with amax * 2^e in [2^13, 2^14): the largest entry sits two binades under the fp16 maximum.
__device__ __forceinline__ float pow2_scale(float amax) {
  if (!(amax > 0.f) || !isfinite(amax)) return 1.f;
  int e;
  frexpf(amax, &e);  // amax = m 2^e, m in [0.5, 1)
  int s = 14 - e;
  s = s < -60 ? -60 : (s > 60 ? 60 : s);
  return ldexpf(1.f, s);
}

// covariance: operand a = sqrt(nw)(x - shift); |a| <= sqrt(max nw) * dev, dev = bound or measured
__global__ void syrk_scale_kernel(GemmScalars* s, const float* __restrict__ dev_bound) {
  const float wmax = __uint_as_float(s->amax_a_bits);
  const float dev = dev_bound ? *dev_bound : __uint_as_float(s->amax_b_bits);
  const float sc = pow2_scale(sqrtf(wmax) * dev);
  s->scale_a = s->scale_b = sc;
  s->inv_scale = (1.f / sc) * (1.f / sc);
}

// sampling: A = L, B = z (pair when injected; Philox normals are exact and unscaled)
__global__ void sample_scale_kernel(GemmScalars* s, int z_pair) {
  const float sa = pow2_scale(__uint_as_float(s->amax_a_bits));
  const float sb = z_pair ? pow2_scale(__uint_as_float(s->amax_b_bits)) : 1.f;
  s->scale_a = sa;
  s->scale_b = sb;
  s->inv_scale = (1.f / sa) * (1.f / sb);
}

// Covariance operand: At[i][k] = scale sqrt(nw_k) (x[k0+k][i] - shift_i), transposed so that the
// sample index is contiguous, then split.  Tile = 32 samples x 128 columns through shared memory.
template <int MODE>
__global__ void __launch_bounds__(256)
syrk_operand_kernel(const float* __restrict__ x, const float* __restrict__ nw,
                    const float* __restrict__ shift, int64_t K, int64_t D, int64_t k0, int64_t Kc,
                    int64_t Dp, const GemmScalars* __restrict__ scalars, void* __restrict__ op_a,
                    void* __restrict__ op_b) {
  __shared__ float tile[32][129];
  const float scale = scalars ? scalars->scale_a : 1.f;
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
    for (int u = 0; u < 4; ++u) tile[r][c4 + u] = v[u] * scale;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < 128 * 8; e += 256) {
    const int il = e / 8, k4 = (e % 8) * 4;
    const int64_t i = it + il;
    if (i >= Dp) continue;
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = tile[k4 + u][il];
    store_split4<MODE>(v, i, kt + k4, Kc, op_a, op_b);
  }
}

// Row-major operand (rows, cols) -> split rows of padded length cols_p (zero padded), from src
// (pitch cols; lower_only keeps c <= r) scaled by *scale_ptr, or, when src == NULL, from Philox
// normals of the global (.., cols) base-normal matrix (unscaled).
template <int MODE>
__global__ void __launch_bounds__(256)
split_rows_kernel(const float* __restrict__ src, int64_t rows, int64_t cols, int64_t cols_p,
                  uint64_t seed, uint64_t draw, uint64_t row_offset, int lower_only,
                  const float* __restrict__ scale_ptr, void* __restrict__ op_a, void* __restrict__ op_b) {
  const float scale = (src && scale_ptr) ? *scale_ptr : 1.f;
  const int nq_row = (int)(cols_p / 4);
  const bool quad_rng = (cols & 3) == 0;
  // one row per block iteration; threads stride over the row's quads (no 64-bit divisions)
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    for (int q = threadIdx.x; q < nq_row; q += blockDim.x) {
      const int64_t c = (int64_t)q * 4;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (src) {
        if (quad_rng && !lower_only && c + 3 < cols &&
            ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(src + r * cols + c));
          v[0] = scale * t.x; v[1] = scale * t.y; v[2] = scale * t.z; v[3] = scale * t.w;
        } else {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (c + u < cols && (!lower_only || c + u <= r)) v[u] = scale * __ldg(src + r * cols + c + u);
        }
      } else {   // Philox normals of the dense Gaussian sampler: column-quad order (element by element here;
                 // the FP16X2 scheme generates its store with philox_store_kernel, this is the TF32X3 path)
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (c + u < cols) v[u] = philox_cq_normal1(seed, draw, row_offset + (uint64_t)r, (uint32_t)(c + u));
      }
      store_split4<MODE>(v, r, c, cols_p, op_a, op_b);
    }
  }
}

// Base normals of the dense Gaussian sampler, generated straight into the blocked K-major store
//   zT[k / 2048][i][k % 2048] = z[k][i]      (k = local sample, i = column, exact fp16)
// with the column-quad Philox order (common.cuh): one call = samples 4Q .. 4Q+3 of one column = 8 contiguous
// bytes of the store, so a thread writes 8 samples (two calls) as one 16-byte store and a warp 512 contiguous
// bytes; one block = one whole 4 KB row of a store block.  No shared memory, no transposition.
// Samples >= rows_valid and columns >= D are written as zeros (finite operands for the padded tiles).
__global__ void __launch_bounds__(256, 8)
philox_store_kernel(uint16_t* __restrict__ zt, int64_t store_rows, int64_t n_blocks, int64_t rows_valid,
                    int64_t D, uint64_t seed, uint64_t draw, uint64_t row_offset, int rows_log2) {
  // work item = (sample block of 2048, group of 2^rows_log2 store rows), dealt round-robin to the CTAs.  Inside
  // an item a thread keeps its 8 samples (two row quads) and walks down the rows: the row quads and the
  // alignment / ragged-end decision are loop invariants, the inner loop is branch-free with a 32-bit row
  // counter and one pointer increment -- the per-item bookkeeping of the flat version cost 46 of 212
  // instructions per 16-byte store (ncu).
  const PhiloxKeys keys = philox_keys(seed ^ kPhiloxColTag);
  const bool aligned = (row_offset & 3) == 0;
  const uint32_t groups = (uint32_t)(store_rows >> rows_log2);
  const uint32_t n_items = (uint32_t)n_blocks * groups;
  const uint32_t rows_item = 1u << rows_log2;
  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const uint32_t blk = item / groups, i0 = (item - blk * groups) << rows_log2;
    const int64_t k = (int64_t)blk * kZBlock + (int64_t)threadIdx.x * 8;       // first local sample of this thread
    uint16_t* dst = zt + ((int64_t)blk * store_rows + i0) * kZBlock + (int64_t)threadIdx.x * 8;
    const uint32_t i_end = (uint32_t)max((int64_t)i0, min((int64_t)(i0 + rows_item), D));  // rows >= D: zeros
    if (aligned && k + 8 <= rows_valid) {
      const uint64_t q = (row_offset + (uint64_t)k) >> 2;
#pragma unroll 2
      for (uint32_t i = i0; i < i_end; ++i, dst += kZBlock) {
        __half2 a0, a1, b0, b1;
        philox_cq_normal4_h2(keys, draw, i, q, a0, a1);
        philox_cq_normal4_h2(keys, draw, i, q + 1, b0, b1);
        uint4 o;
        o.x = *reinterpret_cast<const uint32_t*>(&a0);
        o.y = *reinterpret_cast<const uint32_t*>(&a1);
        o.z = *reinterpret_cast<const uint32_t*>(&b0);
        o.w = *reinterpret_cast<const uint32_t*>(&b1);
        *reinterpret_cast<uint4*>(dst) = o;
      }
    } else {   // unaligned shard offset, the ragged end of the draw, or samples beyond it (zeros)
      for (uint32_t i = i0; i < i_end; ++i, dst += kZBlock) {
        uint16_t h[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float v = (k + u < rows_valid) ? philox_cq_normal1(seed, draw, row_offset + (uint64_t)(k + u), i) : 0.f;
          const __half hv = __float2half_rn(v);
          h[u] = *reinterpret_cast<const uint16_t*>(&hv);
        }
        uint4 o;
        o.x = (uint32_t)h[0] | ((uint32_t)h[1] << 16);
        o.y = (uint32_t)h[2] | ((uint32_t)h[3] << 16);
        o.z = (uint32_t)h[4] | ((uint32_t)h[5] << 16);
        o.w = (uint32_t)h[6] | ((uint32_t)h[7] << 16);
        *reinterpret_cast<uint4*>(dst) = o;
      }
    }
    for (uint32_t i = i_end; i < i0 + rows_item; ++i, dst += kZBlock)
      *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
  }
}

// The same draw as a row-major fp32 matrix (rows, cols): what the store holds, for callers that want to look at
// the base normals (tests, ppi_philox_normal_cq).  Thread = (row quad, column), columns fastest.
__global__ void __launch_bounds__(256)
philox_cq_matrix_kernel(float* __restrict__ out, int64_t rows, int64_t cols, uint64_t seed, uint64_t draw,
                        uint64_t row_offset) {
  const int64_t nq = (rows + 3) / 4 + 1;    // row quads that can intersect [row_offset, row_offset + rows)
  const uint64_t q0 = row_offset >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nq * cols; e += stride) {
    const int64_t qi = e / cols, c = e % cols;
    const Philox4 p = philox4x32_10(q0 + (uint64_t)qi, (draw << 32) | (uint64_t)c, seed ^ kPhiloxColTag);
    float v[4];
    box_muller(p.x, p.y, v[0], v[1]);
    box_muller(p.z, p.w, v[2], v[3]);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t r = (int64_t)((q0 + (uint64_t)qi) * 4 + u) - (int64_t)row_offset;
      if (r >= 0 && r < rows) out[r * cols + c] = v[u];
    }
  }
}

// cov[i][j] (+)= inv_scale * sum_s partial[s][min(i,j)][max(i,j)]
__global__ void syrk_reduce_kernel(const float* __restrict__ partials, int splits, int64_t Dp,
                                   int64_t split_stride, int64_t D, float* __restrict__ cov,
                                   int accumulate, const GemmScalars* __restrict__ scalars) {
  const float inv = scalars ? scalars->inv_scale : 1.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < D * D; e += stride) {
    const int64_t i = e / D, j = e % D;
    const int64_t a = i < j ? i : j, b = i < j ? j : i;
    float s = 0.f;
    for (int sp = 0; sp < splits; ++sp) s += partials[(int64_t)sp * split_stride + a * Dp + b];
    s *= inv;
    cov[e] = accumulate ? cov[e] + s : s;
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
int gemm_scheme();     // capi.cu: PPI_GEMM_TF32X3 / PPI_GEMM_FP16X2 (ppi_set_option)
int gemm_chunk();      // capi.cu: pipeline stages per TMEM accumulation chunk
int sample_chunk();    // capi.cu: the same for the sampling kernels

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

// 2-D K-major operand of `rows` rows; one box = box_rows x one 128-byte swizzle row.
//   OP_TF32X3     fp32 [rows][kp]              box 32 elements  (32 contraction steps)
//   OP_F16_PAIR   fp16 [rows][kp/32][h1|h2]    box 64 elements  (32 contraction steps)
//   OP_F16_SINGLE fp16 [rows][kp]              box 64 elements  (64 contraction steps)
static int make_map(CUtensorMap* map, const void* ptr, int64_t rows, int64_t kp, int box_rows, int mode) {
  EncodeTiledFn fn = get_encode_fn();
  PPI_REQUIRE(fn != nullptr, PPI_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not available");
  const bool f16 = mode != OP_TF32X3;
  const int64_t elems = mode == OP_F16_PAIR ? 2 * kp : kp;
  cuuint64_t dims[2] = {(cuuint64_t)elems, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)elems * (f16 ? 2 : 4)};
  cuuint32_t box[2] = {(cuuint32_t)(f16 ? 64 : 32), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                  (void*)ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PPI_REQUIRE(r == CUDA_SUCCESS, PPI_ERR_DRIVER, "cuTensorMapEncodeTiled failed with CUresult %d",
              (int)r);
  return 0;
}

static int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }
// injected-normal store: sample-major rows of fp16 pairs
static size_t zstore_row_bytes(int64_t D) { return (size_t)round_up(D, 64) * 4; }
// Philox store: K-major and blocked over samples, zT[k / 2048][i][k % 2048] = z[k][i]: blocks of 2048
// samples x round_up(D, 128) rows of exact fp16 (a TMA box of 128 / 256 rows then spans 0.5 / 1 MB instead
// of one row pitch of 2 K bytes per row: measured 30 % faster at K = 2^20, where that pitch is a 2 MB page)
static int64_t zstore_kt(int64_t K) { return round_up(K, kZBlock); }
static int grid_cap(int64_t work, int per_sm) {
  const int64_t cap = (int64_t)kNumSMs * per_sm;
  return (int)(work < 1 ? 1 : (work < cap ? work : cap));
}

template <typename Kern>
static int set_gemm_smem(Kern kern) {
  PPI_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kGemmSmemBytes));
  return 0;
}

// Side stream for the operand pre-passes: staging pass p+1 (HBM-bound) overlaps the GEMM of pass p
// (bound by operand delivery / the tensor pipe) through two operand buffers.  Ordered against the
// caller's stream with events only; one set per device, created on first use.
struct SideStream {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr, op_done[2] = {nullptr, nullptr}, gemm_done[2] = {nullptr, nullptr};
};
static int side_stream(SideStream** out) {
  static SideStream table[64];
  int dev = 0;
  PPI_CUDA_TRY(cudaGetDevice(&dev));
  PPI_REQUIRE(dev >= 0 && dev < 64, PPI_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  SideStream& ss = table[dev];
  if (ss.stream == nullptr) {
    PPI_CUDA_TRY(cudaStreamCreateWithFlags(&ss.stream, cudaStreamNonBlocking));
    PPI_CUDA_TRY(cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming));
    for (int i = 0; i < 2; ++i) {
      PPI_CUDA_TRY(cudaEventCreateWithFlags(&ss.op_done[i], cudaEventDisableTiming));
      PPI_CUDA_TRY(cudaEventCreateWithFlags(&ss.gemm_done[i], cudaEventDisableTiming));
    }
  }
  *out = &ss;
  return 0;
}
int gemm_overlap();  // capi.cu: 1 = stage operands on the side stream
int gemm_pair();     // capi.cu: 1 = CTA-pair (cta_group::2) kernels where they exist
int prefetch_blocks();  // capi.cu: resident CTAs per SM of ppi_philox_prefetch

constexpr int64_t kSyrkKc = 65536;    // samples staged per pass
constexpr int64_t kSampleKc = 65536;  // sample rows staged per pass
constexpr int64_t kKcQuantum = 512;   // Kc granularity (covers every chunk / split rounding below)
constexpr size_t kScalarBytes = 1024; // GemmScalars block at the head of the workspace

struct SyrkPlan {
  int64_t Dp, Kc, tiles_m, tiles_n, ntiles;
  int splits;
  size_t op_bytes, partial_bytes;
};

static SyrkPlan syrk_plan(int64_t K, int64_t D) {
  SyrkPlan pl;
  pl.Dp = round_up(D, BM);
  pl.Kc = round_up(K < kSyrkKc ? K : kSyrkKc, kKcQuantum);
  pl.tiles_m = pl.Dp / BM;
  pl.tiles_n = (pl.Dp + BN - 1) / BN;
  pl.ntiles = 0;
  for (int64_t tm = 0; tm < pl.tiles_m; ++tm) pl.ntiles += pl.tiles_n - (tm * BM) / BN;
  int64_t s = kNumSMs / pl.ntiles;
  const int64_t max_s = pl.Kc / kKcQuantum;
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
  return kScalarBytes + 2 * pl.op_bytes + pl.partial_bytes + 1024;
}

template <bool FP16>
static int wmoments_dense_impl(const float* x, const float* nw, int64_t K, int64_t D,
                               const float* shift, const float* dev_bound, float* cov, int accumulate,
                               void* workspace, cudaStream_t st) {
  constexpr int MODE = FP16 ? OP_F16_PAIR : OP_TF32X3;
  const SyrkPlan pl = syrk_plan(K, D);
  char* wsp = (char*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  GemmScalars* scalars = FP16 ? (GemmScalars*)wsp : nullptr;
  void* op_a = (void*)(wsp + kScalarBytes);
  void* op_b = FP16 ? nullptr : (void*)(wsp + kScalarBytes + pl.op_bytes);
  float* partials = (float*)(wsp + kScalarBytes + 2 * pl.op_bytes);

  if (FP16) {
    // operand scale: sqrt(max nw) * max |x - shift| (given as a bound, or measured with one more pass over x)
    ProfileScope prof(PPI_PROF_SYRK_OPERAND, st);
    PPI_CUDA_TRY(cudaMemsetAsync(scalars, 0, sizeof(GemmScalars), st));
    absmax_kernel<<<grid_cap(ceil_div(K, 1024), 4), 256, 0, st>>>(nw, K, &scalars->amax_a_bits);
    int n = 2;
    if (!dev_bound) {
      absmax_dev_kernel<<<grid_cap(ceil_div(K * D, 4096), 8), 256, 0, st>>>(x, nw, shift, K, D,
                                                                              &scalars->amax_b_bits);
      ++n;
    }
    syrk_scale_kernel<<<1, 1, 0, st>>>(scalars, dev_bound);
    if (int rc = check_launch("syrk scale kernels", n)) return rc;
  }

  // FP16: the second operand array is a second staging buffer (pass p+1 is staged while pass p
  // multiplies); TF32X3 needs both arrays for one pass.
  const bool overlap = FP16 && gemm_overlap() && K > pl.Kc;
  void* bufs[2] = {op_a, FP16 ? (void*)(wsp + kScalarBytes + pl.op_bytes) : op_a};
  CUtensorMap a[2], a2[2], b[2], b2[2];
  for (int i = 0; i < (FP16 ? 2 : 1); ++i) {
    if (int rc = make_map(&a[i], bufs[i], pl.Dp, pl.Kc, BM, MODE)) return rc;
    if (int rc = make_map(&b[i], bufs[i], pl.Dp, pl.Kc, BN, MODE)) return rc;
    a2[i] = a[i];
    b2[i] = b[i];
  }
  if (!FP16) {
    if (int rc = make_map(&a2[0], op_b, pl.Dp, pl.Kc, BM, MODE)) return rc;
    if (int rc = make_map(&b2[0], op_b, pl.Dp, pl.Kc, BN, MODE)) return rc;
  }

  SyrkParams p;
  p.num_kb = (int)(pl.Kc / BK);
  p.chunk_kb = gemm_chunk();
  p.kb_per_split = (int)round_up((p.num_kb + pl.splits - 1) / pl.splits, p.chunk_kb);
  p.splits = pl.splits;
  p.tiles_n = (int)pl.tiles_n;
  p.out = partials;
  p.ld = pl.Dp;
  p.split_stride = pl.Dp * pl.Dp;

  static bool attr_set = false;
  if (!attr_set) {
    if (int rc = set_gemm_smem(syrk_gemm_kernel<FP16 ? SYRK_PAIR : SYRK_TF32X3>)) return rc;
    attr_set = true;
  }
  SideStream* ss = nullptr;
  cudaStream_t st_op = st;
  if (overlap) {
    if (int rc = side_stream(&ss)) return rc;
    st_op = ss->stream;
    PPI_CUDA_TRY(cudaEventRecord(ss->fork, st));
    PPI_CUDA_TRY(cudaStreamWaitEvent(st_op, ss->fork, 0));
  }
  int pass = 0;
  for (int64_t k0 = 0; k0 < K; k0 += pl.Kc, ++pass) {
    const int bi = overlap ? (pass & 1) : 0;
    dim3 og((unsigned)(pl.Kc / 32), (unsigned)ceil_div(pl.Dp, 128));
    {
      if (overlap && pass >= 2) PPI_CUDA_TRY(cudaStreamWaitEvent(st_op, ss->gemm_done[bi], 0));
      ProfileScope prof(PPI_PROF_SYRK_OPERAND, st_op);
      syrk_operand_kernel<MODE><<<og, 256, 0, st_op>>>(x, nw, shift, K, D, k0, pl.Kc, pl.Dp, scalars,
                                                       bufs[bi], op_b);
      if (int rc = check_launch("syrk_operand_kernel")) return rc;
    }
    if (overlap) {
      PPI_CUDA_TRY(cudaEventRecord(ss->op_done[bi], st_op));
      PPI_CUDA_TRY(cudaStreamWaitEvent(st, ss->op_done[bi], 0));
    }
    p.accumulate = pass > 0;
    {
      ProfileScope prof(PPI_PROF_SYRK_GEMM, st);
      syrk_gemm_kernel<FP16 ? SYRK_PAIR : SYRK_TF32X3><<<(unsigned)(pl.ntiles * pl.splits), kGemmThreads, kGemmSmemBytes, st>>>(
          a[bi], a2[bi], b[bi], b2[bi], p);
      if (int rc = check_launch("syrk_gemm_kernel")) return rc;
    }
    if (overlap) PPI_CUDA_TRY(cudaEventRecord(ss->gemm_done[bi], st));
  }
  ProfileScope prof(PPI_PROF_SYRK_REDUCE, st);
  syrk_reduce_kernel<<<grid_cap(ceil_div(D * D, 256), 8), 256, 0, st>>>(partials, pl.splits, pl.Dp,
                                                                         pl.Dp * pl.Dp, D, cov, accumulate,
                                                                         scalars);
  return check_launch("syrk_reduce_kernel");
}

extern "C" int ppi_wmoments_dense(const float* x, const float* nw, int64_t K, int64_t D,
                                  const float* shift, const float* dev_bound, float* cov,
                                  int accumulate, void* workspace, size_t workspace_bytes,
                                  void* stream) {
  PPI_REQUIRE(x && nw && cov && K > 0 && D > 0 && workspace, PPI_ERR_INVALID,
              "ppi_wmoments_dense: bad arguments");
  PPI_REQUIRE(workspace_bytes >= ppi_wmoments_dense_workspace_bytes(K, D), PPI_ERR_WORKSPACE,
              "ppi_wmoments_dense: workspace too small");
  if (gemm_scheme() == PPI_GEMM_FP16X2)
    return wmoments_dense_impl<true>(x, nw, K, D, shift, dev_bound, cov, accumulate, workspace,
                                     (cudaStream_t)stream);
  return wmoments_dense_impl<false>(x, nw, K, D, shift, dev_bound, cov, accumulate, workspace,
                                    (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// Whitened moment match.  The samples of ppi_sample_gaussian_keep are x_k = mu + L z_k, so
//   sum_k nw_k x_k                    = mu + L m,          m = sum_k nw_k z_k
//   sum_k nw_k (x_k - mu')(x_k - mu')^T = L (S - m m^T) L^T,  S = sum_k nw_k z_k z_k^T
// and the contraction runs on the stored base normals instead of x: operand A = nw z (fp16 pair),
// operand B = z, which for Philox normals is ONE exact fp16 -- two partial products instead of
// three, 2 bytes per value read instead of 4, no pass over x at all.
// ---------------------------------------------------------------------------------------------
constexpr float kPhiloxAbsMax = 8.0f;  // |Box-Muller normal| <= sqrt(-2 ln 2^-33) = 6.8 (common.cuh)

// whitened operand scale: A = nw z;  injected normals are stored scaled by zs->scale_b
__global__ void white_scale_kernel(GemmScalars* s, const GemmScalars* zs, int injected) {
  const float wmax = __uint_as_float(s->amax_a_bits);
  const float zmax = injected ? __uint_as_float(zs->amax_b_bits) : kPhiloxAbsMax;
  const float sa = pow2_scale(wmax * zmax);
  const float sb = injected ? zs->scale_b : 1.f;
  s->scale_a = sa;
  s->scale_b = sb;
  s->inv_scale = (1.f / sa) * (1.f / sb);
}

// One block = `tiles` consecutive 32-sample tiles x 128 columns of the store rows of this pass:
//   op_a[i][k] = scale_a nw_k z_ki  (pair, K-major)     op_b[i][k] = z_ki as stored (single / pair)
//   s1[i] += sum_k nw_k z_ki        (fp64, one atomic per column per block)
template <bool ZPAIR>
__global__ void __launch_bounds__(256)
white_operand_kernel(const uint16_t* __restrict__ zrows, const float* __restrict__ nw,
                     int64_t rows_valid, int64_t Dk, int64_t Kc, int64_t Dp, int tiles,
                     const GemmScalars* __restrict__ scalars, void* __restrict__ op_a,
                     void* __restrict__ op_b, double* __restrict__ s1, double* __restrict__ s2, int64_t D) {
  __shared__ float tile[32][129];
  __shared__ float wsc[32];
  const float sa = scalars->scale_a;
  const float unz = 1.f / scalars->scale_b;         // undoes the scale of stored injected normals (1 for Philox)
  const int64_t it = (int64_t)blockIdx.y * 128;     // column offset
  const int64_t row_halfs = ZPAIR ? 2 * Dk : Dk;
  double colsum = 0.0, colsq = 0.0;
  for (int tl = 0; tl < tiles; ++tl) {
    const int64_t kt = ((int64_t)blockIdx.x * tiles + tl) * 32;  // local sample offset in the pass
    if (kt >= Kc) break;
    if (threadIdx.x < 32) {
      const int64_t k = kt + threadIdx.x;
      wsc[threadIdx.x] = k < rows_valid ? __ldg(nw + k) * sa * unz : 0.f;
    }
    for (int e = threadIdx.x; e < 32 * 32; e += 256) {
      const int r = e / 32, c4 = (e % 32) * 4;
      const int64_t k = kt + r, i = it + c4;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (k < rows_valid && i < Dk) {
        const uint16_t* row = zrows + k * row_halfs;
        if (ZPAIR) {
          const uint16_t* b = row + (i >> 5) * 64 + (i & 31);
          const uint2 h1 = __ldg(reinterpret_cast<const uint2*>(b));
          const uint2 h2 = __ldg(reinterpret_cast<const uint2*>(b + 32));
          const float2 a0 = __half22float2(*reinterpret_cast<const __half2*>(&h1.x));
          const float2 a1 = __half22float2(*reinterpret_cast<const __half2*>(&h1.y));
          const float2 b0 = __half22float2(*reinterpret_cast<const __half2*>(&h2.x));
          const float2 b1 = __half22float2(*reinterpret_cast<const __half2*>(&h2.y));
          v[0] = a0.x + b0.x; v[1] = a0.y + b0.y; v[2] = a1.x + b1.x; v[3] = a1.y + b1.y;  // exact
        } else {
          const uint2 h = __ldg(reinterpret_cast<const uint2*>(row + i));
          const float2 a0 = __half22float2(*reinterpret_cast<const __half2*>(&h.x));
          const float2 a1 = __half22float2(*reinterpret_cast<const __half2*>(&h.y));
          v[0] = a0.x; v[1] = a0.y; v[2] = a1.x; v[3] = a1.y;
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) tile[r][c4 + u] = v[u];
    }
    __syncthreads();
    if (threadIdx.x < 128) {
      float cs = 0.f, cq = 0.f;
#pragma unroll 8
      for (int r = 0; r < 32; ++r) {
        const float t = tile[r][threadIdx.x], wt = wsc[r] * t;
        cs += wt;
        cq = fmaf(wt, t, cq);
      }
      colsum += (double)cs;
      colsq += (double)cq;
    }
    for (int e = threadIdx.x; e < 128 * 8; e += 256) {
      const int il = e / 8, k4 = (e % 8) * 4;
      const int64_t i = it + il;
      if (i >= Dp) continue;
      float v[4], wv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        v[u] = tile[k4 + u][il];
        wv[u] = v[u] * wsc[k4 + u];
      }
      store_split4<OP_F16_PAIR>(wv, i, kt + k4, Kc, op_a, nullptr);
      store_split4<ZPAIR ? OP_F16_PAIR : OP_F16_SINGLE>(v, i, kt + k4, Kc, op_b, nullptr);
    }
    __syncthreads();
  }
  if (s1 && threadIdx.x < 128 && it + threadIdx.x < D)
    atomicAdd(s1 + it + threadIdx.x, colsum * (double)(1.f / sa));
  // diagonal of S (see white_mean_bulk_kernel): the tile holds the STORED normals, z = tile / scale_b
  if (s2 && threadIdx.x < 128 && it + threadIdx.x < D)
    atomicAdd(s2 + it + threadIdx.x, colsq * (double)(1.f / sa) * (double)unz);
}

// ---------------------------------------------------------------------------------------------
// Whitened covariance straight from the K-major Philox store: no operand pre-pass.
//   C[m][n] = sum_k (nw_k zT[m][k]) zT[n][k]
// Both operands are TMA boxes of the same fp16 array.  A "transform" warpgroup turns each A box
// (128 columns x 64 samples, exact fp16) into the fp16 pair of scale_a nw_k z in shared memory
// (two 32-sample pair tiles in the canonical K-major 128-byte-swizzle layout) while the previous
// stage multiplies; the B box feeds the MMA as it lands.  m = sum_k nw_k z_k comes from
// white_mean_bulk_kernel (one HBM pass over the 2-byte store).
// CTA = 4 warpgroups: WG0 = TMA producer + MMA issuer, WG1-2 = TMEM drain (fp32 register
// accumulators, as syrk_gemm_kernel), WG3 = transform; register budget moved with setmaxnreg.
// ---------------------------------------------------------------------------------------------
// scale_a nw_k split into two fp16 (wa = rn(v), wb = rn(v - wa)); entries k >= K are zero.
__global__ void __launch_bounds__(256)
white_weights_kernel(const float* __restrict__ nw, int64_t K, int64_t kt, const GemmScalars* __restrict__ scalars,
                     uint16_t* __restrict__ wa, uint16_t* __restrict__ wb) {
  const float sa = scalars->scale_a;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < kt; k += stride) {
    const float v = k < K ? __ldg(nw + k) * sa : 0.f;
    const __half a = __float2half_rn(v);
    const __half b = __float2half_rn(v - __half2float(a));
    wa[k] = *reinterpret_cast<const uint16_t*>(&a);
    wb[k] = *reinterpret_cast<const uint16_t*>(&b);
  }
}

// s1[i] += sum_k nw_k zT[i][k]: block = 8 store rows x 16384 samples, weights loaded once per 8 rows.
constexpr int kMeanRows = 8;
// white_mean_bulk_kernel: s1[i] += sum_k nw_k zT[i][k] and s2[i] += sum_k nw_k zT[i][k]^2 -- the DIAGONAL of S with
// round-to-nearest fp32 products and fp64 sums.  The tensor-core accumulation truncates (a bias of ~2e-6 on sums
// of positive terms), and tr C = sum of D such entries enters the KL as (tr C - D): the finish takes the diagonal
// from here (one HBM pass over the 2-byte store).
// The store is streamed through shared memory by bulk copies (cp.async.bulk, the non-tensor TMA path): the 8 rows
// of a group are adjacent in zT[block][row][2048], so one stage = one 32 KB copy.  Bytes in flight do not depend
// on registers and occupancy (a first version that loaded into registers held 3 CTAs of 80 registers per SM,
// 35 % occupancy, and spent its time in long-scoreboard stalls: ncu 5.0 TB/s, 0.95 ms; this one 6.6 TB/s, 0.65 ms).  CTA = (row group, range
// of sample blocks), two stages, two CTAs per SM; one thread issues the copies, __syncthreads() hands a
// consumed stage back.
constexpr int kMeanStages = 2;
constexpr uint32_t kMeanStageBytes = kMeanRows * (uint32_t)kZBlock * 2u;   // 32 KB
constexpr size_t kMeanBulkSmem = kMeanStages * kMeanStageBytes + 1024;

__global__ void __launch_bounds__(256, 2)
white_mean_bulk_kernel(const uint16_t* __restrict__ zt, int64_t zt_rows, const float* __restrict__ nw, int64_t K,
                       int64_t D, int n_blocks, int ranges, double* __restrict__ s1, double* __restrict__ s2) {
  extern __shared__ uint8_t wm_raw[];
  __shared__ __align__(8) uint64_t full[kMeanStages];
  __shared__ double red[kMeanRows][8];
  __shared__ double red2[kMeanRows][8];
  const uint32_t raw = smem_u32(wm_raw);
  const uint32_t base = (raw + 127u) & ~127u;
  const uint8_t* tiles = wm_raw + (base - raw);
  const int64_t i0 = (int64_t)blockIdx.x * kMeanRows;
  const int b_begin = (int)((int64_t)blockIdx.y * n_blocks / ranges);
  const int b_end = (int)(((int64_t)blockIdx.y + 1) * n_blocks / ranges);
  const int nb = b_end - b_begin;
  const bool w_vec = (reinterpret_cast<uintptr_t>(nw) & 15) == 0;
  auto issue = [&](int it) {     // thread 0: stage it % 2 <- rows i0 .. i0+7 of sample block b_begin + it
    const int s = it % kMeanStages;
    const uint32_t bar = smem_u32(&full[s]);
    mbar_expect_tx(bar, kMeanStageBytes);
    const uint16_t* src = zt + ((int64_t)(b_begin + it) * zt_rows + i0) * kZBlock;
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(base + (uint32_t)s * kMeanStageBytes), "l"(src), "r"(kMeanStageBytes), "r"(bar) : "memory");
  };
  if (threadIdx.x == 0) {
    for (int s = 0; s < kMeanStages; ++s) mbar_init(smem_u32(&full[s]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int it = 0; it < kMeanStages && it < nb; ++it) issue(it);
  }
  __syncthreads();
  float acc[kMeanRows], acc2[kMeanRows];
  double accd[kMeanRows], accd2[kMeanRows];     // fp32 partials are folded into fp64 every 8 stages (64 terms)
#pragma unroll
  for (int r = 0; r < kMeanRows; ++r) { acc[r] = acc2[r] = 0.f; accd[r] = accd2[r] = 0.0; }
  for (int it = 0; it < nb; ++it) {
    const int s = it % kMeanStages;
    const int64_t k = (int64_t)(b_begin + it) * kZBlock + (int64_t)threadIdx.x * 8;
    float w[8];
    if (k + 8 <= K && w_vec) {
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(nw + k));
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(nw + k) + 1);
      w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w; w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
    } else {
#pragma unroll
      for (int u = 0; u < 8; ++u) w[u] = (k + u < K) ? __ldg(nw + k + u) : 0.f;
    }
    if ((it & 7) == 7) {
#pragma unroll
      for (int r = 0; r < kMeanRows; ++r) {
        accd[r] += (double)acc[r]; accd2[r] += (double)acc2[r];
        acc[r] = acc2[r] = 0.f;
      }
    }
    mbar_wait(smem_u32(&full[s]), (uint32_t)(it / kMeanStages) & 1u);
    const uint8_t* tile = tiles + (size_t)s * kMeanStageBytes + (size_t)threadIdx.x * 16;
#pragma unroll
    for (int r = 0; r < kMeanRows; ++r) {
      const uint4 zr = *reinterpret_cast<const uint4*>(tile + (size_t)r * kZBlock * 2);
      const uint32_t zw[4] = {zr.x, zr.y, zr.z, zr.w};
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float2 zf = __half22float2(*reinterpret_cast<const __half2*>(&zw[u]));
        const float wx = zf.x * w[2 * u], wy = zf.y * w[2 * u + 1];
        acc[r] += wx;
        acc[r] += wy;
        acc2[r] = fmaf(wx, zf.x, acc2[r]);
        acc2[r] = fmaf(wy, zf.y, acc2[r]);
      }
    }
    __syncthreads();                                   // stage consumed by every thread
    if (threadIdx.x == 0 && it + kMeanStages < nb) issue(it + kMeanStages);
  }
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
#pragma unroll
  for (int r = 0; r < kMeanRows; ++r) {
    const double v = warp_sum(accd[r] + (double)acc[r]);
    const double v2 = warp_sum(accd2[r] + (double)acc2[r]);
    if (lane == 0) { red[r][wp] = v; red2[r][wp] = v2; }
  }
  __syncthreads();
  if (threadIdx.x < kMeanRows && i0 + threadIdx.x < D) {
    double v = 0.0, v2 = 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q) { v += red[threadIdx.x][q]; v2 += red2[threadIdx.x][q]; }
    atomicAdd(s1 + i0 + threadIdx.x, v);
    if (s2) atomicAdd(s2 + i0 + threadIdx.x, v2);
  }
}

constexpr int kWhiteThreads = 512;
constexpr int kWStages = 3;                               // TMA ring: A source (16 KB) | B (32 KB)
constexpr int kWPairStages = 2;                           // transformed A: 2 pair tiles (32 KB)
constexpr uint32_t kWStageBytes = kABytes + kBBytes;      // 48 KB
constexpr uint32_t kWPairBytes = 2 * kABytes;             // 32 KB
constexpr uint32_t kWPipeBytes = kWStages * kWStageBytes + kWPairStages * kWPairBytes;  // 208 KB
constexpr uint32_t kWhiteSmemBytes = kWPipeBytes + 1024 /*align*/ + 256 /*barriers*/;

struct WhiteParams {
  int num_st;         // 64-sample stages in the store row (kt / 64)
  int chunk_st;       // stages per TMEM accumulation chunk
  int st_per_split;   // stages handled by one CTA
  int splits, tiles_n;
  float* out;         // partials [splits][ld][ld]
  int64_t ld, split_stride;
  const uint16_t* wa; // scale_a nw_k = wa_k + wb_k as two fp16 arrays of kt entries (zero beyond K)
  const uint16_t* wb;
};


__global__ void __launch_bounds__(kWhiteThreads, 1)
syrk_white_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  const WhiteParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t pair_base = base + kWStages * kWStageBytes;
  const uint32_t bar_base = base + kWPipeBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (3 + s); };
  auto pfull_bar = [&](int s) { return bar_base + 8u * (6 + s); };
  auto pempty_bar = [&](int s) { return bar_base + 8u * (8 + s); };
  auto tfull_bar = [&](int b) { return bar_base + 8u * (10 + b); };
  auto tempty_bar = [&](int b) { return bar_base + 8u * (12 + b); };
  uint8_t* smem = smem_raw + (base - raw);
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(smem + kWPipeBytes + 8 * 14);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kWStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < kWPairStages; ++s) { mbar_init(pfull_bar(s), 4); mbar_init(pempty_bar(s), 1); }
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

  // ---- tile / k-range of this CTA (upper-triangular tiles, as syrk_gemm_kernel) ----
  int t = blockIdx.x / p.splits;
  const int split = blockIdx.x % p.splits;
  int tm = 0, tn = 0;
  for (;; ++tm) {
    const int tn_min = (tm * BM) / BN;
    const int cnt = p.tiles_n - tn_min;
    if (t < cnt) { tn = tn_min + t; break; }
    t -= cnt;
  }
  const int st_begin = split * p.st_per_split;
  const int st_end = min(p.num_st, st_begin + p.st_per_split);
  const int nst = max(0, st_end - st_begin);
  const int nchunks = (nst + p.chunk_st - 1) / p.chunk_st;

  if (warp < 4) {
    reg_dec<48>();
    if (warp == 0 && lane == 0 && nst > 0) {
      // ================= TMA producer =================
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
      int stage = 0;
      uint32_t phase = 0;
      for (int st = st_begin; st < st_end; ++st) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t sb = base + stage * kWStageBytes;
        mbar_expect_tx(full_bar(stage), kWStageBytes);
        const int c0 = (st % (int)(kZBlock / 64)) * 64;                 // sample within its 2048-block
        const int rb = (st / (int)(kZBlock / 64)) * (int)p.ld;          // first store row of that block
        tma_load_2d(sb, &map_a, full_bar(stage), c0, rb + tm * BM);
        tma_load_2d(sb + kABytes, &map_b, full_bar(stage), c0, rb + tn * BN);
        if (++stage == kWStages) { stage = 0; phase ^= 1u; }
      }
    } else if (warp == 1 && lane == 0 && nst > 0) {
      // ================= MMA issuer =================
      int stage = 0, ps = 0;
      uint32_t phase = 0, pphase = 0;
      int st = 0;
      for (int c = 0; c < nchunks; ++c) {
        const int buf = c & 1;
        mbar_wait(tempty_bar(buf), (((uint32_t)c >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)buf * BN;
        const int st_chunk_end = min(nst, st + p.chunk_st);
        uint32_t accumulate = 0;
        for (; st < st_chunk_end; ++st) {
          mbar_wait(full_bar(stage), phase);    // B landed
          mbar_wait(pfull_bar(ps), pphase);     // A transformed
          tc_fence_after();
          const uint32_t bt = base + stage * kWStageBytes + kABytes;
          const uint32_t pa = pair_base + ps * kWPairBytes;
          issue_f16_pair_single(tmem_d, pa, bt, 0, accumulate);
          issue_f16_pair_single(tmem_d, pa + kABytes, bt, 1, accumulate);
          tc_commit(empty_bar(stage));
          tc_commit(pempty_bar(ps));
          if (++stage == kWStages) { stage = 0; phase ^= 1u; }
          if (++ps == kWPairStages) { ps = 0; pphase ^= 1u; }
        }
        tc_commit(tfull_bar(buf));
      }
    }
  } else if (warp < 12) {
    reg_inc<176>();
    // ================= drain TMEM chunks into fp32 registers =================
    const int e = warp - 4;
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    const int h = e >> 2;     // column half
    float acc[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) acc[i] = 0.f;
    for (int c = 0; c < nchunks; ++c) {
      const int buf = c & 1;
      mbar_wait(tfull_bar(buf), ((uint32_t)c >> 1) & 1u);
      tc_fence_after();
      drain_add(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN + h * 128), acc);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(buf));
    }
    const int64_t m = (int64_t)tm * BM + q * 32 + lane;
    const int64_t n0 = (int64_t)tn * BN + h * 128;
    if (m < p.ld) {
      float* dst = p.out + (int64_t)split * p.split_stride + m * p.ld + n0;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (n0 + 4 * j < p.ld)
          reinterpret_cast<float4*>(dst)[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
    }
  } else {
    reg_dec<112>();
    // ================= transform: A box (exact fp16 z) -> fp16 pair of (scale_a nw_k) z =================
    // All in packed half arithmetic.  With the weight itself split as wa + wb (fp16 each, white_weights_kernel):
    //   h1 = rn(wa z)            HMUL2
    //   r  = wa z - h1           HFMA2, exact (the rounding error of an fp16 product is an fp16 number)
    //   h2 = rn(wb z + r)        HFMA2
    // so h1 + h2 = (wa + wb) z up to 2^-22 relative, three instructions per two values.
    const int tt = threadIdx.x - 384;   // 0..127
    const int c = tt & 7;               // 16-byte chunk of the 128-byte source row: samples 8c .. 8c+7
    const int r0 = tt >> 3;             // rows r0 + 16 j
    int stage = 0, ps = 0;
    uint32_t phase = 0, pphase = 0;
    // the split weights of the next stage are fetched one stage ahead (an L2 round trip per stage would
    // otherwise sit on the transform -> MMA critical path); the arrays are zero-padded to kt
    uint4 wa_n = make_uint4(0, 0, 0, 0), wb_n = wa_n;
    if (nst > 0) {
      const int64_t k = (int64_t)st_begin * 64 + c * 8;
      wa_n = __ldg(reinterpret_cast<const uint4*>(p.wa + k));
      wb_n = __ldg(reinterpret_cast<const uint4*>(p.wb + k));
    }
    for (int st = st_begin; st < st_end; ++st) {
      const uint32_t wa[4] = {wa_n.x, wa_n.y, wa_n.z, wa_n.w}, wb[4] = {wb_n.x, wb_n.y, wb_n.z, wb_n.w};
      if (st + 1 < st_end) {
        const int64_t k = (int64_t)(st + 1) * 64 + c * 8;
        wa_n = __ldg(reinterpret_cast<const uint4*>(p.wa + k));
        wb_n = __ldg(reinterpret_cast<const uint4*>(p.wb + k));
      }
      mbar_wait(full_bar(stage), phase);
      mbar_wait(pempty_bar(ps), pphase ^ 1u);
      const uint8_t* src = smem + stage * kWStageBytes;
      uint8_t* dst = smem + kWStages * kWStageBytes + ps * kWPairBytes + (c >> 2) * kABytes;
      const int jj = c & 3;
      const bool lo_half = c < 4;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int r = r0 + 16 * j;
        const int x = r & 7;
        const uint4 zr = *reinterpret_cast<const uint4*>(src + r * 128 + ((c ^ x) << 4));
        const uint32_t zw[4] = {zr.x, zr.y, zr.z, zr.w};
        uint32_t h1[4], h2[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const __half2 z2 = *reinterpret_cast<const __half2*>(&zw[u]);
          const __half2 a2 = *reinterpret_cast<const __half2*>(&wa[u]);
          const __half2 b2 = *reinterpret_cast<const __half2*>(&wb[u]);
          const __half2 hi = __hmul2_rn(a2, z2);
          const __half2 rem = __hfma2(a2, z2, __hneg2(hi));
          const __half2 lo = __hfma2(b2, z2, rem);
          h1[u] = *reinterpret_cast<const uint32_t*>(&hi);
          h2[u] = *reinterpret_cast<const uint32_t*>(&lo);
        }
        // lanes c < 4 (first pair tile) store h1 then h2, lanes c >= 4 (second tile, same banks 16 KB further)
        // h2 then h1: the 8 lanes of a quarter warp always hit 8 different 16-byte bank groups
        const uint4 v1 = make_uint4(h1[0], h1[1], h1[2], h1[3]), v2 = make_uint4(h2[0], h2[1], h2[2], h2[3]);
        const int p1 = (jj ^ x) << 4, p2 = ((4 + jj) ^ x) << 4;
        *reinterpret_cast<uint4*>(dst + r * 128 + (lo_half ? p1 : p2)) = lo_half ? v1 : v2;
        *reinterpret_cast<uint4*>(dst + r * 128 + (lo_half ? p2 : p1)) = lo_half ? v2 : v1;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> tcgen05 reads
      __syncwarp();
      if (lane == 0) mbar_arrive(pfull_bar(ps));
      if (++stage == kWStages) { stage = 0; phase ^= 1u; }
      if (++ps == kWPairStages) { ps = 0; pphase ^= 1u; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

static size_t white_op_bytes(const SyrkPlan& pl, int injected) {
  // A pair (4 B) + B single (2 B) or pair (4 B) per (column, sample)
  return (size_t)pl.Dp * pl.Kc * (injected ? 8 : 6);
}

extern "C" size_t ppi_wmoments_white_workspace_bytes(int64_t K, int64_t D, int injected) {
  if (K <= 0 || D <= 0) return 0;
  const SyrkPlan pl = syrk_plan(K, D);
  if (!injected)  // direct kernel: scalars + one fp32 partial per split (splits <= SMs / tiles)
    return kScalarBytes + (size_t)round_up(4 * zstore_kt(K), 1024) +
           (size_t)(kNumSMs / pl.ntiles < 1 ? 1 : kNumSMs / pl.ntiles) * pl.Dp * pl.Dp * sizeof(float) + 2048;
  return kScalarBytes + white_op_bytes(pl, injected) + pl.partial_bytes + 2048;
}

template <bool ZPAIR>
static int wmoments_white_impl(const void* zstore, const float* nw, int64_t K, int64_t D, double* s1, double* s2,
                               float* cov, int accumulate, void* workspace, cudaStream_t st) {
  constexpr int MODE = ZPAIR ? SYRK_PAIR : SYRK_PAIR_SINGLE;
  const SyrkPlan pl = syrk_plan(K, D);
  const int64_t Dk = round_up(D, 64);
  char* wsp = (char*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  GemmScalars* scalars = (GemmScalars*)wsp;
  const GemmScalars* zs = (const GemmScalars*)zstore;
  const char* store_rows = (const char*)zstore + kScalarBytes;
  const size_t store_row_bytes = zstore_row_bytes(D);
  void* op_a = (void*)(wsp + kScalarBytes);
  void* op_b = (void*)(wsp + kScalarBytes + (size_t)pl.Dp * pl.Kc * 4);
  float* partials = (float*)(wsp + kScalarBytes + white_op_bytes(pl, ZPAIR ? 1 : 0));
  partials = (float*)(((uintptr_t)partials + 1023) & ~(uintptr_t)1023);

  {
    ProfileScope prof(PPI_PROF_SYRK_OPERAND, st);
    PPI_CUDA_TRY(cudaMemsetAsync(scalars, 0, sizeof(GemmScalars), st));
    absmax_kernel<<<grid_cap(ceil_div(K, 1024), 4), 256, 0, st>>>(nw, K, &scalars->amax_a_bits);
    white_scale_kernel<<<1, 1, 0, st>>>(scalars, zs, ZPAIR ? 1 : 0);
    if (int rc = check_launch("white scale kernels", 2)) return rc;
  }
  CUtensorMap a, b;
  if (int rc = make_map(&a, op_a, pl.Dp, pl.Kc, BM, OP_F16_PAIR)) return rc;
  if (int rc = make_map(&b, op_b, pl.Dp, pl.Kc, BN, ZPAIR ? OP_F16_PAIR : OP_F16_SINGLE)) return rc;

  SyrkParams p;
  const int unit = ZPAIR ? BK : 2 * BK;                      // contraction steps per pipeline stage
  const int chunk_units = ZPAIR ? gemm_chunk() : gemm_chunk() / 2;  // same 32-step chunk boundaries
  p.num_kb = (int)(pl.Kc / unit);
  p.chunk_kb = chunk_units < 1 ? 1 : chunk_units;
  p.kb_per_split = (int)round_up((p.num_kb + pl.splits - 1) / pl.splits, p.chunk_kb);
  p.splits = pl.splits;
  p.tiles_n = (int)pl.tiles_n;
  p.out = partials;
  p.ld = pl.Dp;
  p.split_stride = pl.Dp * pl.Dp;

  static bool attr_set = false;
  if (!attr_set) {
    if (int rc = set_gemm_smem(syrk_gemm_kernel<MODE>)) return rc;
    attr_set = true;
  }
  constexpr int kTilesPerBlock = 8;
  int pass = 0;
  for (int64_t k0 = 0; k0 < K; k0 += pl.Kc, ++pass) {
    const int64_t rows = (K - k0) < pl.Kc ? (K - k0) : pl.Kc;
    {
      ProfileScope prof(PPI_PROF_SYRK_OPERAND, st);
      dim3 og((unsigned)ceil_div(pl.Kc / 32, kTilesPerBlock), (unsigned)ceil_div(pl.Dp, 128));
      white_operand_kernel<ZPAIR><<<og, 256, 0, st>>>(
          (const uint16_t*)(store_rows + (size_t)k0 * store_row_bytes), nw + k0, rows, Dk, pl.Kc, pl.Dp,
          kTilesPerBlock, scalars, op_a, op_b, s1, s2, D);
      if (int rc = check_launch("white_operand_kernel")) return rc;
    }
    p.accumulate = pass > 0;
    {
      ProfileScope prof(PPI_PROF_SYRK_GEMM, st);
      syrk_gemm_kernel<MODE><<<(unsigned)(pl.ntiles * pl.splits), kGemmThreads, kGemmSmemBytes, st>>>(
          a, a, b, b, p);
      if (int rc = check_launch("syrk_gemm_kernel(white)")) return rc;
    }
  }
  ProfileScope prof(PPI_PROF_SYRK_REDUCE, st);
  syrk_reduce_kernel<<<grid_cap(ceil_div(D * D, 256), 8), 256, 0, st>>>(partials, pl.splits, pl.Dp,
                                                                         pl.Dp * pl.Dp, D, cov, accumulate,
                                                                         scalars);
  return check_launch("syrk_reduce_kernel");
}

// Philox store (K-major): one launch over all K samples, no operand pre-pass.
static int wmoments_white_direct(const void* zstore, const float* nw, int64_t K, int64_t D, double* s1, double* s2,
                                 float* cov, int accumulate, void* workspace, cudaStream_t st) {
  const SyrkPlan pl = syrk_plan(K, D);
  const int64_t kt = zstore_kt(K);
  char* wsp = (char*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  GemmScalars* scalars = (GemmScalars*)wsp;
  uint16_t* wa = (uint16_t*)(wsp + kScalarBytes);
  uint16_t* wb = wa + kt;
  float* partials = (float*)(wsp + kScalarBytes + round_up(4 * kt, 1024));
  const uint16_t* zt = (const uint16_t*)((const char*)zstore + kScalarBytes);
  {
    ProfileScope prof(PPI_PROF_SYRK_OPERAND, st);
    PPI_CUDA_TRY(cudaMemsetAsync(scalars, 0, sizeof(GemmScalars), st));
    absmax_kernel<<<grid_cap(ceil_div(K, 1024), 4), 256, 0, st>>>(nw, K, &scalars->amax_a_bits);
    white_scale_kernel<<<1, 1, 0, st>>>(scalars, (const GemmScalars*)zstore, 0);
    white_weights_kernel<<<grid_cap(ceil_div(kt, 256), 8), 256, 0, st>>>(nw, K, kt, scalars, wa, wb);
    if (int rc = check_launch("white scale kernels", 3)) return rc;
  }
  if (s1) {
    ProfileScope prof(PPI_PROF_WMOMENTS_DIAG, st);
    // samples per CTA: as long as the grid still holds ~8 CTAs per SM (a 16 384-sample CTA lives 6 us: launch and
    // reduction tails cost a quarter of the kernel, ncu: 57 % of the DRAM peak)
    static bool mean_attr = false;
    if (!mean_attr) {
      PPI_CUDA_TRY(cudaFuncSetAttribute(white_mean_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)kMeanBulkSmem));
      mean_attr = true;
    }
    // CTA = (group of 8 store rows, range of 2048-sample blocks): ~7 waves of 2 CTAs per SM
    const int64_t groups = pl.Dp / kMeanRows, n_blocks = kt / kZBlock;
    int64_t ranges = ceil_div((int64_t)kNumSMs * 2 * 7, groups);
    if (ranges > n_blocks) ranges = n_blocks;
    if (ranges < 1) ranges = 1;
    dim3 mg((unsigned)groups, (unsigned)ranges);
    white_mean_bulk_kernel<<<mg, 256, kMeanBulkSmem, st>>>(zt, pl.Dp, nw, K, D, (int)n_blocks, (int)ranges, s1, s2);
    if (int rc = check_launch("white_mean_bulk_kernel")) return rc;
  }
  CUtensorMap a, b;
  if (int rc = make_map(&a, zt, (kt / kZBlock) * pl.Dp, kZBlock, BM, OP_F16_SINGLE)) return rc;
  if (int rc = make_map(&b, zt, (kt / kZBlock) * pl.Dp, kZBlock, BN, OP_F16_SINGLE)) return rc;
  WhiteParams p;
  p.num_st = (int)(kt / 64);
  p.chunk_st = gemm_chunk() / 2 < 1 ? 1 : gemm_chunk() / 2;
  int64_t splits = kNumSMs / pl.ntiles;
  if (splits > p.num_st / p.chunk_st) splits = p.num_st / p.chunk_st;
  if (splits < 1) splits = 1;
  p.splits = (int)splits;
  p.st_per_split = (int)round_up((p.num_st + splits - 1) / splits, p.chunk_st);
  p.tiles_n = (int)pl.tiles_n;
  p.out = partials;
  p.ld = pl.Dp;
  p.split_stride = pl.Dp * pl.Dp;
  p.wa = wa;
  p.wb = wb;
  static bool attr_set = false;
  if (!attr_set) {
    PPI_CUDA_TRY(cudaFuncSetAttribute(syrk_white_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)kWhiteSmemBytes));
    attr_set = true;
  }
  {
    ProfileScope prof(PPI_PROF_SYRK_GEMM, st);
    syrk_white_kernel<<<(unsigned)(pl.ntiles * splits), kWhiteThreads, kWhiteSmemBytes, st>>>(a, b, p);
    if (int rc = check_launch("syrk_white_kernel")) return rc;
  }
  ProfileScope prof(PPI_PROF_SYRK_REDUCE, st);
  syrk_reduce_kernel<<<grid_cap(ceil_div(D * D, 256), 8), 256, 0, st>>>(partials, (int)splits, pl.Dp,
                                                                         pl.Dp * pl.Dp, D, cov, accumulate,
                                                                         scalars);
  return check_launch("syrk_reduce_kernel");
}

extern "C" int ppi_wmoments_white(const void* zstore, int injected, const float* nw, int64_t K,
                                  int64_t D, double* s1, double* s2, float* cov, int accumulate, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(zstore && nw && cov && K > 0 && D > 128 && workspace, PPI_ERR_INVALID,
              "ppi_wmoments_white: bad arguments");
  PPI_REQUIRE(workspace_bytes >= ppi_wmoments_white_workspace_bytes(K, D, injected), PPI_ERR_WORKSPACE,
              "ppi_wmoments_white: workspace too small");
  if (injected)
    return wmoments_white_impl<true>(zstore, nw, K, D, s1, s2, cov, accumulate, workspace, (cudaStream_t)stream);
  return wmoments_white_direct(zstore, nw, K, D, s1, s2, cov, accumulate, workspace, (cudaStream_t)stream);
}

// sample.cu
int ppi_sample_gaussian_small(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                              int64_t K, int D, const float* mu, const float* L, float* x,
                              cudaStream_t st);

static int64_t sample_kc(int64_t K) { return round_up(K < kSampleKc ? K : kSampleKc, BN); }
// staging bytes of one pass: pair / tf32 rows (4 B per value) or, for Philox normals, a mini store of whole
// 2048-sample blocks (2 B per value, round_up(D, 128) rows per block)
static size_t sample_z_bytes(int64_t Kc, int64_t D) {
  const size_t rows = (size_t)Kc * round_up(D, 64) * sizeof(float);
  const size_t store = (size_t)round_up(Kc, kZBlock) * round_up(D, BM) * sizeof(uint16_t);
  return rows > store ? rows : store;
}

extern "C" size_t ppi_sample_gaussian_workspace_bytes(int64_t K, int64_t D) {
  if (K <= 0 || D <= 128) return 0;
  const int64_t Dk = round_up(D, 64);
  return kScalarBytes + 2 * (size_t)D * Dk * sizeof(float) + 2 * sample_z_bytes(sample_kc(K), D) + 1024;
}

// sum of the per-tile partial costs in a fixed order + the function's affine part
__global__ void __launch_bounds__(256)
cost_finish_kernel(const float* __restrict__ part, int64_t K, int tiles, int fn, float D,
                   float* __restrict__ cost) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < K; k += stride) {
    float s = 0.f;
    for (int t = 0; t < tiles; ++t) s += part[k * tiles + t];
    if (fn == PPI_FN_STYBLINSKI) s = fmaf(0.5f, s, 39.16599f * D);
    else s += 10.f * D;
    cost[k] = s;
  }
}

// Philox base normals of local samples [0, rows_valid) of a draw into a blocked K-major store of n_blocks
// 2048-sample blocks (column-quad order; the rest of the last block and the padded rows are zero-filled).
static int generate_store(uint16_t* zt, int64_t n_blocks, int64_t rows_valid, int64_t D, uint64_t seed,
                          uint64_t draw_id, uint64_t row_offset, int blocks_per_sm, cudaStream_t st) {
  const int64_t store_rows = round_up(D, BM);
  ProfileScope prof(PPI_PROF_SAMPLE_OPERAND, st);
  // an SM changes its L1 / shared-memory split only when it is empty: ask for the largest shared-memory
  // carve-out although the generator uses none, so that kernels with 100+ KB of dynamic shared memory
  // (dense64.cu chol_step_kernel, the tcgen05 GEMMs) can become resident BESIDE it on another stream
  static const cudaError_t carve = cudaFuncSetAttribute(
      philox_store_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
  (void)carve;
  // rows per work item: as many as keeps >= 6 items per resident CTA (balance) -- 128 at K = 2^20, 16 at 2^17
  const int64_t ctas = (int64_t)kNumSMs * blocks_per_sm;
  int rows_log2 = 7;
  while (rows_log2 > 3 && n_blocks * (store_rows >> rows_log2) < 6 * ctas) --rows_log2;
  const int64_t n_items = n_blocks * (store_rows >> rows_log2);
  philox_store_kernel<<<(unsigned)(n_items < ctas ? n_items : ctas), 256, 0, st>>>(
      zt, store_rows, n_blocks, rows_valid, D, seed, draw_id, row_offset, rows_log2);
  return check_launch("philox_store_kernel");
}

// FP16: L -> fp16 pair; z -> one exact fp16 (Philox, ZPAIR = false) or an fp16 pair (injected).
// !FP16: TF32X3 (ZPAIR ignored; the low-order z operand is skipped for Philox normals).
// Base-normal store kept for the whitened moment match (ppi_sample_gaussian_keep /
// ppi_wmoments_white): a 1 KB header (GemmScalars of the sampling call: scale_b = power-of-two scale
// of the stored injected normals, amax_b_bits = max |z|) followed by
//   Philox    zT[k / 2048][i][k % 2048] = z[k][i]: blocks of 2048 samples x round_up(D, 128) rows of exact
//             fp16 (2 B per value).  K-major for the contraction over samples (covariance) and read as an
//             MN-major operand by the contraction over columns (sampling): the ONLY copy of the base normals
//   injected  round_up(K, 256) sample rows of round_up(D, 64) / 32 blocks [h1 x 32 | h2 x 32] fp16
//             (4 B per value, scaled by scale_b)
template <bool FP16, bool ZPAIR>
static int sample_gaussian_impl(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                                int64_t K, int64_t D, const float* mu, const float* L, float* x,
                                void* workspace, void* zstore, cudaStream_t st,
                                bool store_ready = false, int cost_fn = -1, float* cost_out = nullptr) {
  // store_ready: the Philox store was generated ahead of time (ppi_philox_prefetch): only the GEMMs run
  constexpr int LMODE = FP16 ? OP_F16_PAIR : OP_TF32X3;
  constexpr int ZMODE = FP16 ? (ZPAIR ? OP_F16_PAIR : OP_F16_SINGLE) : OP_TF32X3;
  constexpr bool kStore = FP16 && !ZPAIR;       // Philox normals: the GEMM reads a blocked K-major store
  constexpr int KS = kStore ? 64 : 32;
  const int64_t Dk = round_up(D, 64), Dp = round_up(D, BM);
  const int64_t Kc = sample_kc(K);
  char* wsp = (char*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  const size_t l_bytes = (size_t)D * Dk * sizeof(float), z_bytes = sample_z_bytes(Kc, D);
  GemmScalars* scalars = FP16 ? (GemmScalars*)wsp : nullptr;
  void* l_a = (void*)(wsp + kScalarBytes);
  void* l_b = (void*)(wsp + kScalarBytes + l_bytes);
  void* z_a = (void*)(wsp + kScalarBytes + 2 * l_bytes);
  void* z_b = (void*)(wsp + kScalarBytes + 2 * l_bytes + z_bytes);
  const bool z_exact = (z == nullptr);  // Philox normals sit on the fp16 (hence TF32) grid (common.cuh box_muller)

  if (FP16) {
    PPI_CUDA_TRY(cudaMemsetAsync(scalars, 0, sizeof(GemmScalars), st));
    absmax_kernel<<<grid_cap(ceil_div(D * D, 1024), 4), 256, 0, st>>>(L, D * D, &scalars->amax_a_bits);
    int n = 2;
    if (ZPAIR) {
      absmax_kernel<<<grid_cap(ceil_div(K * D, 4096), 8), 256, 0, st>>>(z, K * D, &scalars->amax_b_bits);
      ++n;
    }
    sample_scale_kernel<<<1, 1, 0, st>>>(scalars, ZPAIR ? 1 : 0);
    if (int rc = check_launch("sample scale kernels", n)) return rc;
  }
  split_rows_kernel<LMODE><<<grid_cap(D, 8), 256, 0, st>>>(
      L, D, D, Dk, 0, 0, 0, 1, FP16 ? &scalars->scale_a : nullptr, l_a, l_b);
  if (int rc = check_launch("split_rows_kernel(L)")) return rc;

  // injected normals + store: every pass stages its (pair) rows straight into the store (no ping-pong);
  // Philox + store: the whole store is generated up front, the passes read it in place;
  // Philox without a store: every pass generates a mini store of its own samples in the workspace
  char* store_rows = zstore ? (char*)zstore + kScalarBytes : nullptr;
  const size_t store_row_bytes = zstore_row_bytes(D);
  const bool store_pairs = zstore && ZPAIR;
  const bool whole_store = zstore && kStore;
  const bool overlap = FP16 && gemm_overlap() && K > Kc && !store_pairs && !whole_store;
  void* zbufs[2] = {z_a, FP16 ? z_b : z_a};
  CUtensorMap ml, ml2, mz[2], mz2[2];
  if (int rc = make_map(&ml, l_a, D, Dk, BM, LMODE)) return rc;
  ml2 = ml;
  if (kStore) {
    if (whole_store) {
      const int64_t n_blocks = zstore_kt(K) / kZBlock;
      if (!store_ready)
        if (int rc = generate_store((uint16_t*)store_rows, n_blocks, K, D, seed, draw_id, row_offset, 8, st)) return rc;
      if (int rc = make_map(&mz[0], store_rows, n_blocks * Dp, kZBlock, 64, OP_F16_SINGLE)) return rc;
      mz[1] = mz[0];
    } else {
      const int64_t n_blocks = round_up(Kc, kZBlock) / kZBlock;
      for (int i = 0; i < 2; ++i)
        if (int rc = make_map(&mz[i], zbufs[i], n_blocks * Dp, kZBlock, 64, OP_F16_SINGLE)) return rc;
    }
    mz2[0] = mz[0];
    mz2[1] = mz[1];
  } else {
    for (int i = 0; i < (FP16 ? 2 : 1) && !store_pairs; ++i) {
      if (int rc = make_map(&mz[i], zbufs[i], Kc, Dk, BN, ZMODE)) return rc;
      mz2[i] = mz[i];
    }
  }
  if (zstore && FP16)  // header: the scales of this call (read by ppi_wmoments_white)
    PPI_CUDA_TRY(cudaMemcpyAsync(zstore, scalars, sizeof(GemmScalars), cudaMemcpyDeviceToDevice, st));
  if (!FP16) {
    if (int rc = make_map(&ml2, l_b, D, Dk, BM, LMODE)) return rc;
    if (int rc = make_map(&mz2[0], z_b, Kc, Dk, BN, ZMODE)) return rc;
  }

  static bool attr_set = false;
  if (!attr_set) {
    if (int rc = set_gemm_smem(sample_gemm_kernel<FP16, ZPAIR>)) return rc;
    attr_set = true;
  }
  SampleParams p;
  p.tiles_j = (int)ceil_div(D, BM);
  p.units_per_block = (p.tiles_j + 1) / 2;
  p.num_st = (int)(Dk / KS);
  p.chunk_st = kStore ? sample_chunk() / 2 : sample_chunk();  // same 32-step boundaries in every variant
  p.D = D;
  p.mu = mu;
  p.scalars = scalars;
  p.z_exact = z_exact ? 1 : 0;
  // fused cost: per-(sample, j-tile) partials behind the staging buffers of the workspace
  p.cost_part = cost_out ? (float*)(wsp + kScalarBytes + 2 * l_bytes + 2 * z_bytes) : nullptr;
  p.cost_fn = cost_fn;
  p.row0 = 0;
  p.store_k0 = 0;
  p.store_rows = (int)Dp;

  SideStream* ss = nullptr;
  cudaStream_t st_op = st;
  if (overlap) {
    if (int rc = side_stream(&ss)) return rc;
    st_op = ss->stream;
    PPI_CUDA_TRY(cudaEventRecord(ss->fork, st));
    PPI_CUDA_TRY(cudaStreamWaitEvent(st_op, ss->fork, 0));
  }
  int pass = 0;
  for (int64_t k0 = 0; k0 < K; k0 += Kc, ++pass) {
    const int bi = overlap ? (pass & 1) : 0;
    const int64_t rows = (K - k0) < Kc ? (K - k0) : Kc;
    const int64_t rows_p = round_up(rows, BN);
    p.row0 = k0;
    p.store_k0 = whole_store ? k0 : 0;
    if (store_pairs) {
      zbufs[0] = store_rows + (size_t)k0 * store_row_bytes;
      if (int rc = make_map(&mz[0], zbufs[0], rows_p, Dk, BN, ZMODE)) return rc;
      mz2[0] = mz[0];
    }
    if (!whole_store) {
      // injected: rows beyond `rows` keep stale workspace values whose products are never stored;
      // Philox: the mini store is zero-filled beyond `rows`
      if (overlap && pass >= 2) PPI_CUDA_TRY(cudaStreamWaitEvent(st_op, ss->gemm_done[bi], 0));
      if (kStore) {
        if (int rc = generate_store((uint16_t*)zbufs[bi], round_up(rows_p, kZBlock) / kZBlock, rows, D, seed, draw_id,
                                    row_offset + (uint64_t)k0, 8, st_op))
          return rc;
      } else {
        ProfileScope prof(PPI_PROF_SAMPLE_OPERAND, st_op);
        const int64_t nrows = z ? rows : rows_p;
        split_rows_kernel<ZMODE><<<grid_cap(nrows, 8), 256, 0, st_op>>>(
            z ? z + k0 * D : nullptr, nrows, D, Dk, seed, draw_id, row_offset + (uint64_t)k0, 0,
            FP16 ? &scalars->scale_b : nullptr, zbufs[bi], (!FP16 && !z_exact) ? z_b : nullptr);
        if (int rc = check_launch("split_rows_kernel(z)")) return rc;
      }
    }
    if (overlap) {
      PPI_CUDA_TRY(cudaEventRecord(ss->op_done[bi], st_op));
      PPI_CUDA_TRY(cudaStreamWaitEvent(st, ss->op_done[bi], 0));
    }
    p.n_units = (rows_p / BN) * p.units_per_block;
    p.rows = rows;
    p.x = x ? x + k0 * D : nullptr;
    const unsigned g = (unsigned)(p.n_units < kNumSMs ? p.n_units : kNumSMs);
    {
      ProfileScope prof(PPI_PROF_SAMPLE_GEMM, st);
      sample_gemm_kernel<FP16, ZPAIR><<<g, kSampleThreads, kGemmSmemBytes, st>>>(ml, ml2, mz[bi], mz2[bi], p);
      if (int rc = check_launch("sample_gemm_kernel")) return rc;
    }
    if (overlap) PPI_CUDA_TRY(cudaEventRecord(ss->gemm_done[bi], st));
  }
  if (cost_out) {
    ProfileScope prof(PPI_PROF_COST, st);
    cost_finish_kernel<<<grid_cap(ceil_div(K, 256), 8), 256, 0, st>>>(p.cost_part, K, p.tiles_j, cost_fn, (float)D,
                                                                      cost_out);
    if (int rc = check_launch("cost_finish_kernel")) return rc;
  }
  return 0;
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
  if (gemm_scheme() == PPI_GEMM_FP16X2) {
    if (z) return sample_gaussian_impl<true, true>(z, seed, draw_id, row_offset, K, D, mu, L, x, workspace, nullptr, st);
    return sample_gaussian_impl<true, false>(z, seed, draw_id, row_offset, K, D, mu, L, x, workspace, nullptr, st);
  }
  return sample_gaussian_impl<false, true>(z, seed, draw_id, row_offset, K, D, mu, L, x, workspace, nullptr, st);
}

extern "C" size_t ppi_sample_gaussian_cost_workspace_bytes(int64_t K, int64_t D) {
  if (K <= 0 || D <= 128) return 0;
  return ppi_sample_gaussian_workspace_bytes(K, D) + (size_t)K * ceil_div(D, BM) * sizeof(float) + 1024;
}

extern "C" size_t ppi_zstore_bytes(int64_t K, int64_t D, int injected) {
  if (K <= 0 || D <= 128) return 0;
  if (injected) return kScalarBytes + (size_t)round_up(K, BN) * zstore_row_bytes(D) + 1024;
  return kScalarBytes + (size_t)round_up(D, BM) * zstore_kt(K) * 2 + 1024;
}

// The base normals of the dense Gaussian sampler (D > 128) as a row-major fp32 matrix: what
// ppi_sample_gaussian / _keep / _cost draw on the device for (seed, draw_id, row_offset) -- the column-quad
// Philox order of their K-major store, not the row-quad order of ppi_philox_normal.
extern "C" int ppi_philox_normal_cq(float* out, int64_t rows, int64_t cols, uint64_t seed, uint64_t draw_id,
                                    uint64_t row_offset, void* stream) {
  PPI_REQUIRE(out && rows >= 0 && cols > 0, PPI_ERR_INVALID, "ppi_philox_normal_cq: bad arguments");
  if (rows == 0) return 0;
  const int64_t work = ((rows + 3) / 4 + 1) * cols;
  philox_cq_matrix_kernel<<<grid_cap(ceil_div(work, 256), 8), 256, 0, (cudaStream_t)stream>>>(out, rows, cols, seed,
                                                                                              draw_id, row_offset);
  return check_launch("philox_cq_matrix_kernel");
}

// Base normals of a whole Philox draw generated ahead of the sampling call (they depend on (seed, draw,
// row) only, not on the policy): the K-major store, one launch.  Issued on a side stream it runs under the
// HBM- and latency-bound part of the previous update (cost, temperature solve, dense fp64 tail).
extern "C" int ppi_philox_prefetch(uint64_t seed, uint64_t draw_id, uint64_t row_offset, int64_t K,
                                   int64_t D, void* zstore, size_t zstore_bytes, void* stream) {
  PPI_REQUIRE(K > 0 && D > 128 && zstore, PPI_ERR_INVALID, "ppi_philox_prefetch: bad arguments");
  PPI_REQUIRE(((uintptr_t)zstore & 1023) == 0 && zstore_bytes >= ppi_zstore_bytes(K, D, 0), PPI_ERR_WORKSPACE,
              "ppi_philox_prefetch: store misaligned (1 KB) or too small");
  uint16_t* store_rows = (uint16_t*)((char*)zstore + kScalarBytes);
  return generate_store(store_rows, zstore_kt(K) / kZBlock, K, D, seed, draw_id, row_offset, prefetch_blocks(),
                        (cudaStream_t)stream);
}

// ppi_sample_gaussian_keep for a draw whose base normals were prefetched: zstore as written by
// ppi_philox_prefetch for the same (K, D); only the GEMM passes run.
extern "C" int ppi_sample_gaussian_pre(int64_t K, int64_t D, const float* mu, const float* L, float* x,
                                       void* zstore, void* workspace, size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(mu && L && x && K > 0 && D > 128 && zstore, PPI_ERR_INVALID,
              "ppi_sample_gaussian_pre: bad arguments (needs K > 0, D > 128, a prefetched store)");
  PPI_REQUIRE(gemm_scheme() == PPI_GEMM_FP16X2, PPI_ERR_UNSUPPORTED,
              "ppi_sample_gaussian_pre: the base-normal store exists under PPI_GEMM_FP16X2 only");
  PPI_REQUIRE(((uintptr_t)zstore & 1023) == 0, PPI_ERR_WORKSPACE, "ppi_sample_gaussian_pre: store misaligned (1 KB)");
  PPI_REQUIRE(workspace && workspace_bytes >= ppi_sample_gaussian_workspace_bytes(K, D),
              PPI_ERR_WORKSPACE, "ppi_sample_gaussian_pre: workspace too small");
  return sample_gaussian_impl<true, false>(nullptr, 0, 0, 0, K, D, mu, L, x, workspace, zstore,
                                           (cudaStream_t)stream, true);
}

extern "C" int ppi_sample_gaussian_keep(const float* z, uint64_t seed, uint64_t draw_id,
                                        uint64_t row_offset, int64_t K, int64_t D, const float* mu,
                                        const float* L, float* x, void* zstore, size_t zstore_bytes,
                                        void* workspace, size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(mu && L && x && K > 0 && D > 128 && zstore, PPI_ERR_INVALID,
              "ppi_sample_gaussian_keep: bad arguments (needs K > 0, D > 128, a store)");
  PPI_REQUIRE(gemm_scheme() == PPI_GEMM_FP16X2, PPI_ERR_UNSUPPORTED,
              "ppi_sample_gaussian_keep: the base-normal store exists under PPI_GEMM_FP16X2 only");
  PPI_REQUIRE(((uintptr_t)zstore & 1023) == 0 && zstore_bytes >= ppi_zstore_bytes(K, D, z != nullptr),
              PPI_ERR_WORKSPACE, "ppi_sample_gaussian_keep: store misaligned (1 KB) or too small");
  PPI_REQUIRE(workspace && workspace_bytes >= ppi_sample_gaussian_workspace_bytes(K, D),
              PPI_ERR_WORKSPACE, "ppi_sample_gaussian_keep: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (z) return sample_gaussian_impl<true, true>(z, seed, draw_id, row_offset, K, D, mu, L, x, workspace, zstore, st);
  return sample_gaussian_impl<true, false>(z, seed, draw_id, row_offset, K, D, mu, L, x, workspace, zstore, st);
}

// ppi_sample_gaussian_keep with a separable test function evaluated in the sampling epilogue
// (ppi/algorithms.py:85-86 fused: the samples are costed where they are produced).  x may be NULL: the
// samples are then never written (the whitened update contracts the stored base normals; a caller that
// wants x later re-draws it from the same (seed, draw_id, row_offset) with ppi_sample_gaussian).
extern "C" int ppi_sample_gaussian_cost(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                                        int64_t K, int64_t D, const float* mu, const float* L, float* x,
                                        void* zstore, size_t zstore_bytes, int store_ready, int fn, float* cost,
                                        void* workspace, size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(mu && L && cost && K > 0 && D > 128 && zstore, PPI_ERR_INVALID,
              "ppi_sample_gaussian_cost: bad arguments (needs K > 0, D > 128, a store, a cost buffer)");
  PPI_REQUIRE(fn == PPI_FN_RASTRIGIN || fn == PPI_FN_STYBLINSKI, PPI_ERR_UNSUPPORTED,
              "ppi_sample_gaussian_cost: only the separable test functions (Rastrigin, Styblinski) are fused");
  PPI_REQUIRE(gemm_scheme() == PPI_GEMM_FP16X2, PPI_ERR_UNSUPPORTED,
              "ppi_sample_gaussian_cost: the base-normal store exists under PPI_GEMM_FP16X2 only");
  PPI_REQUIRE(((uintptr_t)zstore & 1023) == 0 && zstore_bytes >= ppi_zstore_bytes(K, D, z != nullptr),
              PPI_ERR_WORKSPACE, "ppi_sample_gaussian_cost: store misaligned (1 KB) or too small");
  PPI_REQUIRE(workspace && workspace_bytes >= ppi_sample_gaussian_cost_workspace_bytes(K, D), PPI_ERR_WORKSPACE,
              "ppi_sample_gaussian_cost: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (z)
    return sample_gaussian_impl<true, true>(z, seed, draw_id, row_offset, K, D, mu, L, x, workspace, zstore, st,
                                            false, fn, cost);
  return sample_gaussian_impl<true, false>(z, seed, draw_id, row_offset, K, D, mu, L, x, workspace, zstore, st,
                                           store_ready != 0, fn, cost);
}
