// Gibbs-posterior weights.
//   NaN filter                 Base.filter                ppi/algorithms.py:57-68
//   cost normalisation         (c-min)/(max-min+EPS)      ppi/algorithms.py:184,510,541,597 (:472 MPPI)
//   REPS KL dual               Reps.update                ppi/algorithms.py:183-215
//   ESS target                 Essps.update               ppi/algorithms.py:596-615
//   lower bound                Lbps.log_w                 ppi/algorithms.py:539-567
//   fixed temperature          Ais.log_w / MppiBase       ppi/algorithms.py:509-511,471-473
//   elites                     Cem.update / iCem.update   ppi/algorithms.py:117-120,149-154
//   softmax head of m_projection                          ppi/policies.py:84-87
//
// One cooperative kernel per update: every block keeps its slice of the costs L1-resident and
// the whole grid iterates the 1-D solve together.  Each objective evaluation is a max-shifted
// log-sum-exp pass (the normalised costs have min 0, so the shift is exact and free): fp32
// exponentials, fp64 accumulators, warp-shuffle + one grid barrier per evaluation.  The scalar
// solver (bracket expansion + Illinois regula falsi on a monotone function) runs redundantly and
// deterministically in thread 0 of every block, so no broadcast is needed.
//
// Instead of minimising |ESS - N| or the lower bound with Brent (what SciPy does for the
// reference), the kernel finds the root of the monotone residual / of the analytic derivative.
// The reference's Brent solutions agree with those roots to ~1e-8.
#include "common.cuh"

#include <mutex>
#include <optional>

namespace ppi {

constexpr int kSolveThreads = 256;
constexpr int kMaxSolveBlocks = 4 * kNumSMs;  // cooperative grid: up to 4 co-resident blocks per SM
constexpr int NP = 8;               // doubles per block partial
constexpr int kModeUserLogW = 100;  // internal: input holds log-weights, not costs
constexpr double kAlphaLower = 1e-5, kAlphaUpper = 5e2;  // ppi/algorithms.py:36-37 (REPS only)
// A proposal that moves the iterate by less than this (relative) ends the solve.  The objective is evaluated with
// fp32 exponentials, so the residual carries noise of ~1e-9 and the root is defined to ~1e-10 relative: a 1e-11
// criterion is never met and the solver dithers for 3 - 8 further evaluations before the bracket-width test stops
// it (CPU emulation of this state machine, round 2).  The step that meets 1e-8 is a Newton / secant step: the
// iterate it produces is accurate to ~1e-9, four orders inside the 1e-5 parity bound on alpha.
constexpr double kStepTol = 1e-8;

struct SolveWorkspace {
  unsigned int barrier;
  unsigned int pad[31];
  double partials[2][kMaxSolveBlocks][NP];
};

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// All blocks are co-resident (cooperative launch).  `target` is carried per thread.
__device__ __forceinline__ void grid_sync(unsigned int* bar, unsigned int nblocks,
                                          unsigned int& target) {
  target += nblocks;
  __syncthreads();
  if (nblocks == 1) return;  // single block (K <= 4096, the MPC shapes): the block barrier is the grid barrier
  {
    // the whole grid is ONE thread-block cluster (K <= 65536): the hardware cluster barrier with
    // release / acquire semantics replaces the atomic counter and its spin loop
    unsigned int cluster_ctas;
    asm volatile("mov.u32 %0, %%cluster_nctaid.x;" : "=r"(cluster_ctas));
    if (cluster_ctas == nblocks) {
      asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
      return;
    }
  }
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    while (ld_acquire_u32(bar) < target) {
    }
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct Sums {
  double s0, s1, sc, scc, s1c, sccc, s1cc;  // sum e, e^2, e c, e c^2, e^2 c, e c^3, e^2 c^2
};

// Grid-wide reduction of NV doubles.  vals[] in: this thread's partial; out: grid total in
// every thread of every block, bitwise identical across blocks.
template <int NV>
__device__ __forceinline__ void grid_reduce_sum(double (&vals)[NV], SolveWorkspace* ws, int& buf,
                                                unsigned int& bar_target, double* s_scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = kSolveThreads / 32;
#pragma unroll
  for (int v = 0; v < NV; ++v) vals[v] = warp_sum_d(vals[v]);
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int v = 0; v < NV; ++v) s_scratch[warp * NV + v] = vals[v];
  }
  __syncthreads();
  if (gridDim.x == 1) {
    // single block (K <= 4096, the MPC shapes): the warp partials in shared memory are the grid's;
    // no round trip through global memory, every thread adds them in the same order
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      double t = 0.0;
#pragma unroll
      for (int w = 0; w < NW; ++w) t += s_scratch[w * NV + v];
      vals[v] = t;
    }
    __syncthreads();
    return;
  }
  if (threadIdx.x < NV) {
    double t = 0.0;
    for (int w = 0; w < NW; ++w) t += s_scratch[w * NV + threadIdx.x];
    ws->partials[buf][blockIdx.x][threadIdx.x] = t;
  }
  grid_sync(&ws->barrier, gridDim.x, bar_target);
  if (warp == 0) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      double t = 0.0;
      for (int g = lane; g < (int)gridDim.x; g += 32) t += ws->partials[buf][g][v];
      t = warp_sum_d(t);
      if (lane == 0) s_scratch[v] = t;
    }
  }
  __syncthreads();
#pragma unroll
  for (int v = 0; v < NV; ++v) vals[v] = s_scratch[v];
  buf ^= 1;
  __syncthreads();
}

struct CostView {
  const float* c;     // raw costs or log-weights
  int64_t begin, end; // this block's slice
  double shift;       // cost: min; logw: max
  double scale;       // cost: 1/(max-min) (0 when flat) or 1 for the MPPI shift; logw: 1
  bool all_nan;
  bool user_logw;
};

// normalised cost c_ >= 0 (or max - logw >= 0), and whether the element takes part
__device__ __forceinline__ bool elem(const CostView& v, int64_t k, double& cn) {
  const float c = v.c[k];
  if (v.user_logw) {
    cn = (c != c) ? INFINITY : (v.shift - (double)c);
    return true;
  }
  if (v.all_nan) {
    cn = 0.0;
    return true;
  }
  if (c != c) return false;
  cn = ((double)c - v.shift) * v.scale;
  return true;
}

__device__ __forceinline__ Sums evaluate(const CostView& v, double alpha, SolveWorkspace* ws,
                                         int& buf, unsigned int& bar_target, double* s_scratch) {
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  for (int64_t k = v.begin + threadIdx.x; k < v.end; k += kSolveThreads) {
    double cn;
    if (!elem(v, k, cn)) continue;
    const float t = (float)(-alpha * cn);
    const double e = (double)expf(t);
    if (e > 0.0) {
      const double e2 = e * e, ec = e * cn, e2c = e2 * cn;
      acc[0] += e;
      acc[1] += e2;
      acc[2] += ec;
      acc[3] += ec * cn;
      acc[4] += e2c;
      acc[5] += ec * cn * cn;
      acc[6] += e2c * cn;
    }
  }
  grid_reduce_sum<7>(acc, ws, buf, bar_target, s_scratch);
  return Sums{acc[0], acc[1], acc[2], acc[3], acc[4], acc[5], acc[6]};
}

// residual whose root is the sought alpha; `dir` = +1 when it increases with alpha
__device__ __forceinline__ double residual(int mode, const Sums& s, double alpha, double hyper,
                                           double n_valid) {
  switch (mode) {
    case PPI_TEMP_ESS_TARGET:  // log ESS(alpha) - log N   (decreasing)
      return 2.0 * log(s.s0) - log(s.s1) - log(hyper);
    case PPI_TEMP_REPS_KL:     // KL(nw || uniform) - epsilon   (increasing)
      return -alpha * (s.sc / s.s0) - log(s.s0) + log(n_valid) - hyper;
    default: {                 // d/dalpha [E_nw c + lambda / sqrt(ESS)]   (negative at 0)
      const double lambda = sqrt((1.0 - hyper) / hyper);
      const double ep = s.sc / s.s0, var = s.scc / s.s0 - ep * ep;
      const double eq = s.s1c / s.s1, ess = s.s0 * s.s0 / s.s1;
      return -var + lambda * (ep - eq) / sqrt(ess);
    }
  }
}

// d residual / d alpha from the same sums (moments of c under nw ~ e and under nw^2 ~ e^2): lets the
// solver take Newton steps, 2-3 evaluations from a bracket to 1e-11 instead of ~8 regula-falsi steps
__device__ __forceinline__ double residual_slope(int mode, const Sums& s, double alpha, double hyper) {
  const double ep = s.sc / s.s0, eq = s.s1c / s.s1;
  const double var = s.scc / s.s0 - ep * ep;
  switch (mode) {
    case PPI_TEMP_ESS_TARGET:
      return 2.0 * (eq - ep);
    case PPI_TEMP_REPS_KL:
      return alpha * var;
    default: {
      const double lambda = sqrt((1.0 - hyper) / hyper);
      const double m2 = s.scc / s.s0, m3 = s.sccc / s.s0;
      const double k3 = m3 - 3.0 * ep * m2 + 2.0 * ep * ep * ep;  // third central moment under nw
      const double var_q = s.s1cc / s.s1 - eq * eq;
      const double ess = s.s0 * s.s0 / s.s1;
      return k3 + lambda * (-var + 2.0 * var_q + (ep - eq) * (ep - eq)) / sqrt(ess);
    }
  }
}

__device__ __forceinline__ double objective_value(int mode, const Sums& s, double alpha,
                                                  double hyper, double n_valid) {
  const double ess = s.s0 * s.s0 / s.s1;
  switch (mode) {
    case PPI_TEMP_ESS_TARGET:
      return fabs(ess - hyper);                                         // algorithms.py:604
    case PPI_TEMP_REPS_KL:
      return alpha > 0.0 ? hyper / alpha + log(s.s0 / n_valid) / alpha : 0.0;  // :190
    case PPI_TEMP_LOWER_BOUND:
      return s.sc / s.s0 + sqrt((1.0 - hyper) / hyper) / sqrt(ess);     // :553-555
    default:
      return 0.0;
  }
}

struct SolverState {
  int phase;  // 0: evaluate lo, 1: expand hi, 2: refine, 3: done
  double lo, flo, hi, fhi, x, x_prev;
  int last_side;  // -1 lo replaced, +1 hi replaced
  double hi_cap;
  double guess;   // where a Gaussian cost distribution would put the root (NaN: none)
  int left_steps; // consecutive trial points left of the root while expanding
  int status;
};

__global__ void __launch_bounds__(kSolveThreads)
solve_temperature_kernel(const float* __restrict__ in, int64_t K, int mode, double hyper,
                         float* __restrict__ nw, float* __restrict__ log_w,
                         ppi_weight_stats* __restrict__ stats, SolveWorkspace* ws) {
  __shared__ double s_scratch[(kSolveThreads / 32) * NP];
  __shared__ double s_alpha;
  __shared__ int s_done;
  __shared__ SolverState st;

  unsigned int bar_target = 0;
  int buf = 0;
  const bool user = (mode == kModeUserLogW);
  const int64_t chunk = ceil_div(ceil_div(K, (int64_t)gridDim.x), kSolveThreads) * kSolveThreads;
  const int64_t begin = min(K, (int64_t)blockIdx.x * chunk);
  const int64_t end = min(K, begin + chunk);

  // ---------------- phase A: NaN mask, min / max / first arg-extreme, mean, std ---------------
  float vmin = INFINITY, vmax = -INFINITY;
  int64_t imin = INT64_MAX, imax = INT64_MAX;
  double cnt = 0.0, sum = 0.0, sumsq = 0.0;
  for (int64_t k = begin + threadIdx.x; k < end; k += kSolveThreads) {
    const float c = in[k];
    if (c != c) continue;
    if (c < vmin) { vmin = c; imin = k; }
    if (c > vmax) { vmax = c; imax = k; }
    cnt += 1.0;
    sum += (double)c;
    sumsq += (double)c * (double)c;
  }
  {
    // min/max with first-occurrence index: reduce (value, index) pairs lexicographically
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float omin = __shfl_xor_sync(0xffffffffu, vmin, o);
      const int64_t oimin = __shfl_xor_sync(0xffffffffu, imin, o);
      if (omin < vmin || (omin == vmin && oimin < imin)) { vmin = omin; imin = oimin; }
      const float omax = __shfl_xor_sync(0xffffffffu, vmax, o);
      const int64_t oimax = __shfl_xor_sync(0xffffffffu, imax, o);
      if (omax > vmax || (omax == vmax && oimax < imax)) { vmax = omax; imax = oimax; }
    }
    __shared__ float s_min[kSolveThreads / 32], s_max[kSolveThreads / 32];
    __shared__ int64_t s_imin[kSolveThreads / 32], s_imax[kSolveThreads / 32];
    if (lane == 0) { s_min[warp] = vmin; s_imin[warp] = imin; s_max[warp] = vmax; s_imax[warp] = imax; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < kSolveThreads / 32; ++w) {
        if (s_min[w] < vmin || (s_min[w] == vmin && s_imin[w] < imin)) { vmin = s_min[w]; imin = s_imin[w]; }
        if (s_max[w] > vmax || (s_max[w] == vmax && s_imax[w] < imax)) { vmax = s_max[w]; imax = s_imax[w]; }
      }
      double* p = ws->partials[buf][blockIdx.x];
      p[0] = (double)vmin; p[1] = (double)imin; p[2] = (double)vmax; p[3] = (double)imax;
    }
  }
  double a3[3] = {cnt, sum, sumsq};
  // the (min,max) partial rows ride on the same barrier as the sums: write them to the *other*
  // half of the slot (entries 4..7 are used by the sum reduction, 0..3 by min/max)
  {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int v = 0; v < 3; ++v) a3[v] = warp_sum_d(a3[v]);
    __syncthreads();
    if (lane == 0) for (int v = 0; v < 3; ++v) s_scratch[warp * 3 + v] = a3[v];
    __syncthreads();
    if (threadIdx.x < 3) {
      double t = 0.0;
      for (int w = 0; w < kSolveThreads / 32; ++w) t += s_scratch[w * 3 + threadIdx.x];
      ws->partials[buf][blockIdx.x][4 + threadIdx.x] = t;
    }
  }
  grid_sync(&ws->barrier, gridDim.x, bar_target);
  double gmin = INFINITY, gmax = -INFINITY, gimin = 9e18, gimax = 9e18, n_valid = 0.0, gsum = 0.0,
         gsumsq = 0.0;
  for (int g = 0; g < (int)gridDim.x; ++g) {  // every thread redundantly: G <= 148, L2 hits
    const double* p = ws->partials[buf][g];
    if (p[0] < gmin || (p[0] == gmin && p[1] < gimin)) { gmin = p[0]; gimin = p[1]; }
    if (p[2] > gmax || (p[2] == gmax && p[3] < gimax)) { gmax = p[2]; gimax = p[3]; }
    n_valid += p[4]; gsum += p[5]; gsumsq += p[6];
  }
  buf ^= 1;

  const bool all_nan = (n_valid == 0.0);
  if (all_nan) { gmin = 0.0; gmax = 0.0; gimin = 0.0; gimax = 0.0; n_valid = (double)K; gsum = 0.0; gsumsq = 0.0; }
  CostView view;
  view.c = in; view.begin = begin; view.end = end; view.all_nan = all_nan && !user; view.user_logw = user;
  if (user) {
    view.shift = gmax; view.scale = 1.0;
  } else if (mode == PPI_TEMP_FIXED_SHIFTED) {
    view.shift = gmin; view.scale = 1.0;
  } else {
    view.shift = gmin;
    const double range = gmax - gmin;
    view.scale = range > 0.0 ? 1.0 / range : 0.0;  // (c-min)/(0+EPS) = 0 when all costs are equal
  }

  // ---------------- phase B: 1-D solve ----------------------------------------------------------
  const bool fixed = user || mode == PPI_TEMP_FIXED_NORMALISED || mode == PPI_TEMP_FIXED_SHIFTED;
  int iterations = 0;
  if (threadIdx.x == 0) {
    st.phase = fixed ? 3 : 0;
    st.status = 0;
    st.last_side = 0;
    st.lo = (mode == PPI_TEMP_REPS_KL) ? kAlphaLower : 0.0;
    st.hi = 1.0;
    st.hi_cap = (mode == PPI_TEMP_REPS_KL) ? kAlphaUpper : 1.0737418240e9;  // 2^30
    st.x = user ? 1.0 : (fixed ? hyper : st.lo);
    st.x_prev = -1.0;
    st.guess = nan("");
    st.left_steps = 0;
    s_alpha = st.x;
    s_done = fixed ? 1 : 0;
    if (!fixed) {
      // First trial point from the statistics phase A already holds.  If the normalised costs were Gaussian
      // with standard deviation sd, the weights exp(-alpha c) would give ESS = n exp(-(alpha sd)^2),
      // KL(nw || uniform) = (alpha sd)^2 / 2, and the lower bound's derivative would vanish where
      // u exp(u^2 / 2) = sd sqrt(n) / lambda, u = alpha sd.  Real costs are not Gaussian; the point only has to
      // land within a small factor of the root, from there every evaluation is a Newton step.  (Starting at 1
      // and following tangents cost 9 - 28 objective evaluations, tools/solve_iterations.py.)
      const double mean_c = gsum / n_valid;
      const double sd = sqrt(fmax(0.0, gsumsq / n_valid - mean_c * mean_c)) * view.scale;
      if (sd > 0.0 && isfinite(sd)) {
        if (mode == PPI_TEMP_ESS_TARGET) {
          if (hyper < n_valid) st.guess = sqrt(log(n_valid / hyper)) / sd;
        } else if (mode == PPI_TEMP_REPS_KL) {
          st.guess = sqrt(2.0 * hyper) / sd;
        } else {
          const double lr = log(sd) + 0.5 * log(n_valid) - 0.5 * log((1.0 - hyper) / hyper);  // log of the right side
          double u = lr > 0.5 ? sqrt(2.0 * lr) : exp(fmin(lr, 0.5));
          for (int it = 0; it < 8; ++it) u = fmax(1e-6, u - (log(u) + 0.5 * u * u - lr) / (1.0 / u + u));
          st.guess = u / sd;
        }
        if (!(st.guess > st.lo && st.guess < st.hi_cap)) st.guess = nan("");
      }
      // the value at alpha = 0 is known without a pass over the costs: ESS(0) = n, and the bound's derivative
      // at 0 is minus the variance of the normalised costs
      if (mode == PPI_TEMP_ESS_TARGET || mode == PPI_TEMP_LOWER_BOUND) {
        st.flo = (mode == PPI_TEMP_ESS_TARGET) ? log(hyper) - log(n_valid) : -(sd * sd);
        if (!(st.flo < 0.0)) {   // already at / beyond the root at the lower end
          st.x = st.lo; st.phase = 3; st.status = 1; s_done = 1;
        } else {
          if (isfinite(st.guess)) st.hi = st.guess;
          st.phase = 1; st.x = st.hi;
        }
        s_alpha = st.x;
      }
    }
  }
  __syncthreads();
  // sign convention: work with r(alpha) * dir so that the function is negative at lo
  const double dir = (mode == PPI_TEMP_ESS_TARGET) ? -1.0 : 1.0;
  while (!s_done) {
    const double alpha = s_alpha;
    const Sums s = evaluate(view, alpha, ws, buf, bar_target, s_scratch);
    ++iterations;
    if (threadIdx.x == 0) {
      const double f = dir * residual(mode, s, alpha, hyper, n_valid);
      // Newton proposal from this point (f increases through the root: a usable slope is positive)
      const double df = dir * residual_slope(mode, s, alpha, hyper);
      const double newton = (df > 0.0 && isfinite(df)) ? alpha - f / df : nan("");
      if (st.phase == 0) {
        st.flo = f;
        if (!(f < 0.0)) {  // already at / beyond the root at the lower end (or NaN)
          st.x = st.lo; st.phase = 3; st.status = 1;
        } else {
          // first upper point: the tangent's root when there is one (kept within [1e-3, 4]), else 1
          if (isfinite(st.guess)) st.hi = st.guess;
          else if (newton > st.lo && isfinite(newton)) st.hi = fmin(fmin(fmax(newton, 1e-3), 4.0), st.hi_cap);
          st.phase = 1; st.x = st.hi;
        }
      } else if (st.phase == 1) {
        if (f >= 0.0) {
          st.fhi = f; st.phase = 2;
          if (f == 0.0) { st.x = st.hi; st.phase = 3; }
        } else if (!(f < 0.0)) {  // NaN
          st.x = st.lo; st.phase = 3; st.status = 2;
        } else {
          st.lo = st.hi; st.flo = f;
          if (st.hi >= st.hi_cap) { st.x = st.hi_cap; st.phase = 3; st.status = 1; }
          else {
            // still left of the root: follow the tangent (at most 64x, at least the old doubling when
            // there is no slope); a tangent step that no longer moves the iterate is the root
            double nxt = st.hi * 2.0;
            ++st.left_steps;
            if (newton > st.hi && isfinite(newton)) {
              if (newton - st.hi <= kStepTol * st.hi) { st.x = newton; st.phase = 3; }
              // the bound's derivative approaches its root like -C / alpha^2 when the costs have mass at their
              // minimum: tangent steps then only multiply alpha by 1.5.  From the second consecutive miss on the
              // step is doubled each time, until a point right of the root closes the bracket.
              double step = newton - st.hi;
              if (mode == PPI_TEMP_LOWER_BOUND && st.left_steps >= 2) step = ldexp(step, min(st.left_steps - 1, 6));
              nxt = fmin(st.hi + step, st.hi * 64.0);
            }
            if (st.phase != 3) { st.hi = fmin(nxt, st.hi_cap); st.x = st.hi; }
          }
        }
      } else {  // refine
        if (f == 0.0) { st.lo = st.hi = st.x; }
        else if (f > 0.0) {
          st.hi = st.x; st.fhi = f;
          if (st.last_side == 1) st.flo *= 0.5;
          st.last_side = 1;
        } else {
          st.lo = st.x; st.flo = f;
          if (st.last_side == -1) st.fhi *= 0.5;
          st.last_side = -1;
        }
      }
      if (st.phase == 2) {
        const double width = st.hi - st.lo;
        if (width <= 1e-10 * fabs(st.hi) || iterations >= 100) {
          st.x = 0.5 * (st.lo + st.hi);
          st.phase = 3;
          if (iterations >= 100) st.status = 2;
        } else {
          double x = st.hi - st.fhi * width / (st.fhi - st.flo);
          if (!(x > st.lo && x < st.hi)) x = 0.5 * (st.lo + st.hi);
          if (newton > st.lo && newton < st.hi) x = newton;   // tangent from the point just evaluated
          // the iterate has stopped moving: the secant point is the root to fp64 noise level
          if (fabs(x - st.x_prev) <= kStepTol * fabs(x)) st.phase = 3;
          st.x_prev = x;
          st.x = x;
        }
      }
      s_alpha = st.x;
      s_done = (st.phase == 3);
    }
    __syncthreads();
  }

  // ---------------- phase C: final evaluation, normalised weights, stats --------------------------
  const double alpha = s_alpha;
  const Sums s = evaluate(view, alpha, ws, buf, bar_target, s_scratch);
  const double inv_s0 = 1.0 / s.s0;
  for (int64_t k = begin + threadIdx.x; k < end; k += kSolveThreads) {
    double cn;
    float w = 0.f, lw = -INFINITY;
    if (elem(view, k, cn)) {
      const float t = (float)(-alpha * cn);
      w = (float)((double)expf(t) * inv_s0);
      lw = user ? in[k] : t;
    }
    nw[k] = w;
    if (log_w) log_w[k] = lw;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    ppi_weight_stats o;
    o.alpha = user ? 1.0 : alpha;
    const double log_s0 = log(s.s0);
    o.lse = user ? gmax + log_s0 : log_s0;
    o.ess = s.s0 * s.s0 / s.s1;
    o.weighted_cost = s.sc * inv_s0;
    // sum nw log nw = sum nw (t - log s0) = -alpha E[c_] - log s0   (user: t = -(max - lw))
    o.weight_entropy = -(user ? 1.0 : alpha) * s.sc * inv_s0 - log_s0;
    o.cost_min = user ? 0.0 : gmin;
    o.cost_max = user ? 0.0 : gmax;
    const double mean = gsum / n_valid;
    o.cost_mean = user ? 0.0 : mean;
    o.cost_std = user ? 0.0 : sqrt(fmax(0.0, gsumsq / n_valid - mean * mean));
    o.objective = fixed ? 0.0 : objective_value(mode, s, alpha, hyper, n_valid);
    o.n_valid = user ? K : (int64_t)n_valid;
    // np.argmax(log_w): first index of the smallest cost (alpha > 0) / of the largest log-weight
    if (user) o.argmax = (int64_t)gimax;
    else if (alpha > 0.0) o.argmax = (int64_t)gimin;
    else if (alpha < 0.0) o.argmax = (int64_t)gimax;
    else o.argmax = all_nan ? 0 : -1;  // all log-weights equal: first surviving index (scan below)
    o.all_nan = (all_nan && !user) ? 1 : 0;
    o.iterations = iterations;
    o.status = st.status;
    o.pad = 0;
    if (o.argmax < 0) {
      int64_t first = 0;
      while (first < K && in[first] != in[first]) ++first;
      o.argmax = first < K ? first : 0;
    }
    *stats = o;
  }
}

// ------------------------------------------------------------------------------------------
// CEM elites: bitonic sort of (orderable cost, index) keys
// ------------------------------------------------------------------------------------------
constexpr int kSortChunk = 2048;
constexpr int kSortThreads = 1024;

__device__ __forceinline__ uint32_t orderable(float c) {
  if (c != c) return 0xFFFFFFFFu;
  if (c == 0.f) c = 0.f;  // -0 -> +0
  const uint32_t u = __float_as_uint(c);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void make_keys_kernel(const float* __restrict__ c, int64_t K, int64_t N,
                                 const ppi_weight_stats* __restrict__ stats,
                                 uint64_t* __restrict__ keys) {
  const bool all_nan = stats->all_nan != 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
    uint64_t key = ~0ull;
    if (i < K) {
      const uint32_t o = all_nan ? orderable(0.f) : orderable(c[i]);
      key = ((uint64_t)o << 32) | (uint64_t)(uint32_t)i;
    }
    keys[i] = key;
  }
}

__device__ __forceinline__ void cmp_swap(uint64_t& a, uint64_t& b, bool ascending) {
  if ((a > b) == ascending) { const uint64_t t = a; a = b; b = t; }
}

// sorts each kSortChunk block completely (k = 2 .. kSortChunk)
__global__ void __launch_bounds__(kSortThreads) bitonic_local_sort_kernel(uint64_t* keys) {
  __shared__ uint64_t s[kSortChunk];
  const int64_t base = (int64_t)blockIdx.x * kSortChunk;
  for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) s[i] = keys[base + i];
  __syncthreads();
  for (int k = 2; k <= kSortChunk; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      const int p = threadIdx.x;
      const int i = 2 * j * (p / j) + (p % j);
      const bool asc = (((base + i) & k) == 0);
      cmp_swap(s[i], s[i + j], asc);
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) keys[base + i] = s[i];
}

__global__ void bitonic_global_step_kernel(uint64_t* keys, int64_t N, int64_t j, int64_t k) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < N / 2; p += stride) {
    const int64_t i = 2 * j * (p / j) + (p % j);
    const bool asc = ((i & k) == 0);
    uint64_t a = keys[i], b = keys[i + j];
    if ((a > b) == asc) { keys[i] = b; keys[i + j] = a; }
  }
}

// finishes stage k: j = kSortChunk/2 .. 1 inside shared memory
__global__ void __launch_bounds__(kSortThreads) bitonic_local_merge_kernel(uint64_t* keys, int64_t k) {
  __shared__ uint64_t s[kSortChunk];
  const int64_t base = (int64_t)blockIdx.x * kSortChunk;
  for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) s[i] = keys[base + i];
  __syncthreads();
  for (int j = kSortChunk >> 1; j > 0; j >>= 1) {
    const int p = threadIdx.x;
    const int i = 2 * j * (p / j) + (p % j);
    const bool asc = (((base + i) & k) == 0);
    cmp_swap(s[i], s[i + j], asc);
    __syncthreads();
  }
  for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) keys[base + i] = s[i];
}

__global__ void cem_finalize_kernel(const uint64_t* __restrict__ keys, int64_t K, int64_t n_elites,
                                    int64_t* __restrict__ sorted_idx, float* __restrict__ nw,
                                    ppi_weight_stats* __restrict__ stats) {
  const int64_t n_valid = stats->n_valid;
  const int64_t k_eff = n_elites < n_valid ? n_elites : n_valid;
  const float w = 1.0f / (float)k_eff;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < K; r += stride) {
    const int64_t idx = (int64_t)(keys[r] & 0xFFFFFFFFull);
    sorted_idx[r] = idx;
    nw[idx] = r < k_eff ? w : 0.f;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    stats->alpha = 0.0;
    stats->lse = log((double)k_eff);
    stats->ess = (double)k_eff;
    stats->weighted_cost = 0.0;
    stats->weight_entropy = -log((double)k_eff);
    stats->objective = 0.0;
    stats->argmax = (int64_t)(keys[0] & 0xFFFFFFFFull);
  }
}

static int64_t next_pow2(int64_t v) {
  int64_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

static int launch_solve(const float* in, int64_t K, int mode, double hyper, float* nw, float* log_w,
                        ppi_weight_stats* stats, void* workspace, size_t workspace_bytes,
                        cudaStream_t st, bool bracket = true) {
  PPI_REQUIRE(workspace_bytes >= sizeof(SolveWorkspace), PPI_ERR_WORKSPACE,
              "solve workspace too small: %zu < %zu", workspace_bytes, sizeof(SolveWorkspace));
  SolveWorkspace* ws = (SolveWorkspace*)workspace;
  PPI_CUDA_TRY(cudaMemsetAsync(&ws->barrier, 0, sizeof(unsigned int), st));
  static int blocks_per_sm = 0;  // co-residency limit of the cooperative launch
  if (blocks_per_sm == 0) {
    int occ = 0;
    PPI_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, solve_temperature_kernel, kSolveThreads, 0));
    blocks_per_sm = occ < 1 ? 1 : (occ > 4 ? 4 : occ);
  }
  void* args[] = {(void*)&in, (void*)&K, (void*)&mode, (void*)&hyper, (void*)&nw, (void*)&log_w,
                  (void*)&stats, (void*)&ws};
  std::optional<ProfileScope> prof;   // (bracket = false: the allocation-free entry point creates no events)
  if (bracket) prof.emplace(PPI_PROF_SOLVE, st);
  if (K <= 65536) {
    // up to 16 CTAs of 4096 costs each as one cluster (non-portable size, supported on B200; the
    // per-element fp64 work binds one SM to ~0.1 us per 256 costs, so more CTAs beat larger slices):
    // co-scheduled by the hardware, synchronised by the cluster barrier; one CTA needs no barrier at all
    static bool np_set = false;
    if (!np_set) {
      PPI_CUDA_TRY(cudaFuncSetAttribute(solve_temperature_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      np_set = true;
    }
    const unsigned gc = (unsigned)ceil_div(K, 4096);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(gc);
    cfg.blockDim = dim3(kSolveThreads);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = gc;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = gc > 1 ? 1 : 0;
    if (cudaLaunchKernelExC(&cfg, (const void*)solve_temperature_kernel, args) == cudaSuccess) {
      count_launch(1);
      return 0;
    }
    (void)cudaGetLastError();  // fall back to the cooperative grid below
  }
  int64_t g = ceil_div(K, 4096);
  if (g < 1) g = 1;
  if (g > (int64_t)blocks_per_sm * kNumSMs) g = (int64_t)blocks_per_sm * kNumSMs;
  PPI_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)solve_temperature_kernel, dim3((unsigned)g),
                                           dim3(kSolveThreads), args, 0, st));
  count_launch(1);
  return 0;
}

}  // namespace ppi

using namespace ppi;

extern "C" size_t ppi_solve_temperature_workspace_bytes(int64_t K) {
  (void)K;
  return sizeof(SolveWorkspace);
}

extern "C" int ppi_solve_temperature(const float* costs, int64_t K, int mode, double hyper, float* nw,
                                     float* log_w, ppi_weight_stats* stats, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(costs && nw && stats && workspace && K > 0, PPI_ERR_INVALID,
              "ppi_solve_temperature: bad arguments");
  PPI_REQUIRE(mode >= PPI_TEMP_FIXED_NORMALISED && mode <= PPI_TEMP_LOWER_BOUND, PPI_ERR_INVALID,
              "unknown temperature mode %d", mode);
  if (mode == PPI_TEMP_ESS_TARGET) PPI_REQUIRE(hyper > 0.0, PPI_ERR_INVALID, "n_elites must be > 0");
  if (mode == PPI_TEMP_LOWER_BOUND)
    PPI_REQUIRE(hyper > 0.0 && hyper < 1.0, PPI_ERR_INVALID, "delta must be in (0,1)");
  if (mode == PPI_TEMP_REPS_KL) PPI_REQUIRE(hyper > 0.0, PPI_ERR_INVALID, "epsilon must be > 0");
  return launch_solve(costs, K, mode, hyper, nw, log_w, stats, workspace, workspace_bytes,
                      (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// ppi_solve_temperature for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc): the solver's scratch (grid-barrier counter + per-block partial sums, 74 KB) lives in
// a file-scope device array that every GPU holding the library keeps until exit.  In host code the array's NAME
// is a host-side stand-in, not a device address: its address is resolved once per device with
// cudaGetSymbolAddress (that device current) and the resolved pointer is what the kernel gets.
// ---------------------------------------------------------------------------------------------
namespace ppi {
static __device__ __align__(256) unsigned char g_solve_scratch[sizeof(SolveWorkspace)];
constexpr int kMaxScratchDevices = 64;

static int solve_scratch_address(void** out) {
  static std::mutex mu;
  static void* resolved[kMaxScratchDevices] = {};
  int dev = -1;
  PPI_CUDA_TRY(cudaGetDevice(&dev));
  PPI_REQUIRE(dev >= 0 && dev < kMaxScratchDevices, PPI_ERR_UNSUPPORTED,
              "ppi_solve_temperature_static: device ordinal %d outside [0, %d)", dev, kMaxScratchDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (resolved[dev] == nullptr) {
    void* p = nullptr;
    PPI_CUDA_TRY(cudaGetSymbolAddress(&p, g_solve_scratch));
    PPI_REQUIRE(p != nullptr, PPI_ERR_DRIVER, "ppi_solve_temperature_static: scratch symbol did not resolve");
    resolved[dev] = p;
  }
  *out = resolved[dev];
  return 0;
}
}  // namespace ppi

extern "C" int ppi_solve_temperature_static(const float* costs, int64_t K, int mode, double hyper, float* nw,
                                            float* log_w, ppi_weight_stats* stats, void* stream) {
  PPI_REQUIRE(costs && nw && stats && K > 0, PPI_ERR_INVALID, "ppi_solve_temperature_static: bad arguments");
  PPI_REQUIRE(K <= (int64_t)PPI_SOLVE_STATIC_MAX_K, PPI_ERR_UNSUPPORTED,
              "ppi_solve_temperature_static: K = %lld exceeds the maximum of %lld costs of the allocation-free "
              "variant; use ppi_solve_temperature with a caller-owned workspace",
              (long long)K, (long long)PPI_SOLVE_STATIC_MAX_K);
  PPI_REQUIRE(mode >= PPI_TEMP_FIXED_NORMALISED && mode <= PPI_TEMP_LOWER_BOUND, PPI_ERR_INVALID,
              "unknown temperature mode %d", mode);
  if (mode == PPI_TEMP_ESS_TARGET) PPI_REQUIRE(hyper > 0.0, PPI_ERR_INVALID, "n_elites must be > 0");
  if (mode == PPI_TEMP_LOWER_BOUND)
    PPI_REQUIRE(hyper > 0.0 && hyper < 1.0, PPI_ERR_INVALID, "delta must be in (0,1)");
  if (mode == PPI_TEMP_REPS_KL) PPI_REQUIRE(hyper > 0.0, PPI_ERR_INVALID, "epsilon must be > 0");
  void* scratch = nullptr;
  if (int rc = solve_scratch_address(&scratch)) return rc;
  // the barrier counter is cleared on the stream by launch_solve; the partial sums are written before they are
  // read inside every evaluation, so the previous call's contents do not matter
  return launch_solve(costs, K, mode, hyper, nw, log_w, stats, scratch, sizeof(SolveWorkspace),
                      (cudaStream_t)stream, /*bracket=*/false);
}

extern "C" int ppi_normalise_log_weights(const float* log_w, int64_t K, float* nw,
                                         ppi_weight_stats* stats, void* workspace,
                                         size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(log_w && nw && stats && workspace && K > 0, PPI_ERR_INVALID,
              "ppi_normalise_log_weights: bad arguments");
  return launch_solve(log_w, K, kModeUserLogW, 1.0, nw, nullptr, stats, workspace, workspace_bytes,
                      (cudaStream_t)stream);
}

extern "C" size_t ppi_cem_workspace_bytes(int64_t K) {
  int64_t n = next_pow2(K < kSortChunk ? kSortChunk : K);
  return sizeof(SolveWorkspace) + (size_t)n * sizeof(uint64_t);
}

extern "C" int ppi_cem_weights(const float* costs, int64_t K, int64_t n_elites, int64_t* sorted_idx,
                               float* nw, ppi_weight_stats* stats, void* workspace,
                               size_t workspace_bytes, void* stream) {
  PPI_REQUIRE(costs && sorted_idx && nw && stats && workspace && K > 0 && n_elites > 0,
              PPI_ERR_INVALID, "ppi_cem_weights: bad arguments");
  PPI_REQUIRE(K < (1ll << 32), PPI_ERR_UNSUPPORTED, "ppi_cem_weights: K must be < 2^32");
  PPI_REQUIRE(workspace_bytes >= ppi_cem_workspace_bytes(K), PPI_ERR_WORKSPACE,
              "cem workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  // cost statistics + NaN accounting (alpha = 0 pass); nw is overwritten below
  if (int rc = launch_solve(costs, K, PPI_TEMP_FIXED_NORMALISED, 0.0, nw, nullptr, stats, workspace,
                            sizeof(SolveWorkspace), st))
    return rc;
  uint64_t* keys = (uint64_t*)((char*)workspace + sizeof(SolveWorkspace));
  const int64_t N = next_pow2(K < kSortChunk ? kSortChunk : K);
  const int gk = (int)(ceil_div(N, 256) < kNumSMs * 8 ? ceil_div(N, 256) : kNumSMs * 8);
  make_keys_kernel<<<gk, 256, 0, st>>>(costs, K, N, stats, keys);
  const int nchunks = (int)(N / kSortChunk);
  bitonic_local_sort_kernel<<<nchunks, kSortThreads, 0, st>>>(keys);
  int launches = 2;
  for (int64_t k = 2 * kSortChunk; k <= N; k <<= 1) {
    for (int64_t j = k >> 1; j >= kSortChunk; j >>= 1) {
      const int gg = (int)(ceil_div(N / 2, 256) < kNumSMs * 8 ? ceil_div(N / 2, 256) : kNumSMs * 8);
      bitonic_global_step_kernel<<<gg, 256, 0, st>>>(keys, N, j, k);
      ++launches;
    }
    bitonic_local_merge_kernel<<<nchunks, kSortThreads, 0, st>>>(keys, k);
    ++launches;
  }
  const int gf = (int)(ceil_div(K, 256) < kNumSMs * 8 ? ceil_div(K, 256) : kNumSMs * 8);
  cem_finalize_kernel<<<gf, 256, 0, st>>>(keys, K, n_elites, sorted_idx, nw, stats);
  return check_launch("cem kernels", launches + 1);
}
