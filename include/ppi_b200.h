/*
 * ppi_b200.h -- C ABI of the B200-native posterior-policy-iteration (PPI) update path.
 *
 * The reference (JoeMWatson/monte-carlo-posterior-policy-iteration) is pure Python and has no
 * FFI; its boundary is three duck-typed Python protocols (sampler / policy / optimizer).  This
 * header declares what a ctypes binding under those protocols calls.  Every entry point cites the
 * reference interface (file:line relative to the reference root) it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - matrices are dense row-major, fp32, unless typed otherwise;
 *   - `stream` is a cudaStream_t passed as void*; all calls are asynchronous on it;
 *   - the return value is 0 on success, a cudaError_t (>0) for CUDA failures or a negative
 *     PPI_ERR_* code; ppi_last_error() returns a thread-local message;
 *   - the library owns no device memory: workspaces are passed in, sizes come from the matching
 *     *_workspace_bytes query;
 *   - numerical failure (non positive-definite matrix, empty bracket) is reported through device
 *     status words, never through the return value.
 *   - There is NO CPU implementation behind any of these functions.
 */
#ifndef PPI_B200_H
#define PPI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PPI_B200_VERSION 100

#define PPI_ERR_INVALID (-1)     /* bad argument                                    */
#define PPI_ERR_UNSUPPORTED (-2) /* shape outside what the kernels cover            */
#define PPI_ERR_WORKSPACE (-3)   /* workspace too small                             */
#define PPI_ERR_DRIVER (-4)      /* cuTensorMapEncodeTiled unavailable / failed     */

int ppi_version(void);
const char* ppi_last_error(void);

/* Process-wide options (new; no reference counterpart).
 * PPI_OPT_GEMM_SCHEME selects how the tcgen05 kernels (ppi_wmoments_dense, ppi_sample_gaussian)
 * rebuild fp32-accurate products from exact tensor-core operands:
 *   PPI_GEMM_FP16X2  v 2^e = h1 + h2, h1 = fp16(v 2^e), h2 = fp16(rest), one power-of-two scale per
 *                    operand; h1 g1 + h1 g2 + h2 g1 as kind::f16 MMAs (error <= 3 2^-22 |a||b| for
 *                    entries within 2^-28 of the operand's largest; 4 bytes per value).  Default.
 *   PPI_GEMM_TF32X3  hi = tf32(v), lo = tf32(v - hi); hi hi + hi lo + lo hi as kind::tf32 MMAs
 *                    (error ~2^-21 |a||b| at any magnitude; 8 bytes per value, twice the MMA time).
 * PPI_OPT_GEMM_CHUNK = 32-deep k-blocks accumulated inside the tensor core before the partial sum is
 * promoted to fp32 registers (tcgen05 accumulation truncates: ~3e-8 relative per MMA instruction) in the
 * moment kernels.  Even, default 8.  Every promotion drains 128 KB of TMEM, which takes as long as the MMAs of
 * an 8-block chunk; measured on B200 (K = 65536, D = 2048, error relative to the largest entry):
 *   chunk 8: S 2.0e-6 | 16: 3.5e-6 | 32: 7.0e-6 | 64: 1.4e-5   (the diagonal, where the truncation is a
 *   coherent bias, is replaced by the fp64-folded ppi_wmoments_white s2 before the dense tail).
 * PPI_OPT_SAMPLE_CHUNK: the same for the sampling kernels (x = mu + L z).  Even, default 16:
 *   chunk 8: x 3.9e-7, 7.8 ms per 2^20 x 2048 draw | 16: 7.7e-7, 6.9 ms | 32: 1.5e-6, 6.7 ms | 64: 2.9e-6.
 *   The error of x feeds every later quantity (covariance at C2 after three updates: 1.8e-6 / 5.5e-6 with
 *   chunks 8 / 32), so the default keeps a factor of four to the 1e-5 parity bound.
 * PPI_OPT_GEMM_OVERLAP != 0: the FP16X2 kernels stage the operands of pass p+1 on an internal side
 * stream while pass p multiplies (two staging buffers, event-ordered against the caller's stream;
 * the call stays asynchronous and stream-ordered for the caller).  Default 0: measured gain on
 * B200 is ~1 % of a step because both kernels contend for L2 / HBM.
 * PPI_OPT_GEMM_PAIR != 0: kernels that have a CTA-pair form (tcgen05 cta_group::2, clusters of
 * two CTAs sharing the B operand) use it; 0 (default) selects the single-CTA kernels.
 * Results are identical up to fp32 summation order.
 * PPI_OPT_PREFETCH_BLOCKS = resident CTAs per SM of ppi_philox_prefetch, 1..6 (default 6): fewer leave
 * room for the kernels of the stream it runs beside. */
enum { PPI_OPT_GEMM_SCHEME = 0, PPI_OPT_GEMM_CHUNK = 1, PPI_OPT_GEMM_OVERLAP = 2, PPI_OPT_GEMM_PAIR = 3,
       PPI_OPT_PREFETCH_BLOCKS = 4, PPI_OPT_SAMPLE_CHUNK = 5 };
enum { PPI_GEMM_TF32X3 = 0, PPI_GEMM_FP16X2 = 1 };
int ppi_set_option(int key, int value);
int ppi_get_option(int key);

/* ------------------------------------------------------------------------------------------
 * a3  base normals: np.random.randn(n, d)                         ppi/samplers.py:18-20
 * Philox4x32-10 + Box-Muller.  Element (r, c) of the (rows, cols) output is a pure function of
 * (seed, draw_id, row_offset + r, c): K-sharded ranks reproduce the 1-GPU matrix bit for bit.
 * The normals are emitted as IEEE half-precision values held in fp32 (11 significant bits,
 * relative spacing 2^-11) so that they are exact tensor-core operands for ppi_sample_gaussian;
 * every sampler below that draws on the device sees exactly these numbers.
 * ---------------------------------------------------------------------------------------- */
int ppi_philox_normal(float* out, int64_t rows, int64_t cols, uint64_t seed, uint64_t draw_id,
                      uint64_t row_offset, void* stream);
/* The same draw WITHOUT the rounding to the fp16 grid: full fp32 Box-Muller values of the same Philox
 * words (|difference| <= 2^-12 |z|).  These are not exact tensor-core operands: a policy that samples
 * from them goes through the injected-normal path (fp16 pair operands, three partial products).  Used by
 * the distribution tests and by `bench.py --normals fp32`, which prices the quantisation. */
int ppi_philox_normal_full(float* out, int64_t rows, int64_t cols, uint64_t seed, uint64_t draw_id,
                           uint64_t row_offset, void* stream);

/* ------------------------------------------------------------------------------------------
 * f2  quasi-Monte-Carlo base normals                              ppi/samplers.py:25-44
 * out (rows, cols) fp32 = sqrt(2) erfinv(2 shrinkage x - 1), x the points first_index ..
 * first_index + rows - 1 of a scrambled Sobol sequence in `cols` dimensions, bit-identical to
 * scipy.stats.qmc.Sobol(d, scramble=True).random(): sv (cols, bits) and shift (cols) are the
 * engine's direction numbers and digital shift after scrambling (uint32, bits = 30 in SciPy),
 * x_i = 2^-bits (shift ^ XOR_{b set in i ^ (i >> 1)} sv[:, b]).  Every row is a pure function of its
 * index: K-sharded ranks pass their row offset in first_index.
 * ---------------------------------------------------------------------------------------- */
int ppi_sobol_normal(float* out, int64_t rows, int64_t cols, const uint32_t* sv, const uint32_t* shift,
                     int bits, uint64_t first_index, double shrinkage, void* stream);

/* f2  cubature sigma points sqrt(d) [I; -I], out (2 d, d) fp32       ppi/samplers.py:47-53 */
int ppi_cubature_points(float* out, int d, void* stream);

/* ------------------------------------------------------------------------------------------
 * a22 kernel Gram matrix + Cholesky, one thread block, fp64 inside   ppi/policies.py:923-932,
 * 935-974,1026-1037,1087-1091 (Gram) and :312,692-694 (np.linalg.cholesky).
 * kind: 0 squared-exponential, 1 Matern12, 2 Matern32, 3 Matern52, 4 periodic, 5 white noise.
 * t1 (n1), t2 (n2) are fp64 time stamps.  gram64/gram32 are (n1, n2).  If chol64/chol32 are
 * non-NULL (needs n1 == n2 <= 160) the lower Cholesky factor is written too; *status = 0 when
 * positive definite, else 1 + index of the failing pivot.
 * ---------------------------------------------------------------------------------------- */
enum { PPI_KERNEL_SE = 0, PPI_KERNEL_MATERN12 = 1, PPI_KERNEL_MATERN32 = 2,
       PPI_KERNEL_MATERN52 = 3, PPI_KERNEL_PERIODIC = 4, PPI_KERNEL_WHITE = 5 };
int ppi_kernel_gram_chol(const double* t1, int n1, const double* t2, int n2, int kind, double sigma,
                         double lengthscale, double period, double* gram64, float* gram32,
                         double* chol64, float* chol32, int* status, void* stream);

/* ------------------------------------------------------------------------------------------
 * f1  MPC time shift of a kernel prior onto the next window, one launch, fp64
 *     BaseKernel.update_timesteps, ppi/policies.py:761-791 (caller mpc/mpc.py:64-65).
 *   gain = P^-1 (P - Sigma) P^-T;  cross = k(t_new, t_old)
 *   mean_new = clip(cross P^-1 mean + mean_fn, lower, upper) - mean_fn          (n_new, d)
 *   cov_new  = k(t_new, t_new) - anneal cross gain cross^T + jitter I          (n_new, n_new)
 *   chol_new = lower Cholesky factor of cov_new; *status = 0, else 1 + index of the failing pivot
 * prior (P) / prior_inv are the cached Gram matrix of the old window and its inverse (:747-753,
 * 766-768), cov_in (Sigma) and mean (n_old, d) the current moments, all fp64 row-major on the
 * device; jitter = eps * sigma (:785).  lower/upper (d) may both be NULL (no limiter).
 * n_new <= 160.  workspace: ppi_kernel_shift_workspace_bytes(n_old, n_new) bytes.
 * ---------------------------------------------------------------------------------------- */
size_t ppi_kernel_shift_workspace_bytes(int n_old, int n_new);
int ppi_kernel_shift(int kind, double sigma, double lengthscale, double period, const double* t_old,
                     int n_old, const double* t_new, int n_new, const double* prior,
                     const double* prior_inv, const double* cov_in, const double* mean, int d,
                     const double* mean_fn, const double* lower, const double* upper, double anneal,
                     double jitter, double* mean_new, double* cov_new, double* chol_new,
                     double* status, void* workspace, void* stream);

/* np.linalg.cholesky of a small fp64 matrix (n <= 160), one thread block; a may alias l64.
 * ppi/policies.py:312,321,481,819                                                          */
int ppi_cholesky_small(const double* a, int n, double* l64, float* l32, int* status, void* stream);

/* ------------------------------------------------------------------------------------------
 * Sampling.  z == NULL selects on-device Philox normals (seed, draw_id, row_offset as above,
 * the (K, rows*cols) base-normal matrix of ppi/policies.py:361-365); otherwise z holds injected
 * standard normals with the same layout.  lower/upper (per output channel) may both be NULL
 * (no Limiter, ppi/policies.py:65-75,284-285).
 * ---------------------------------------------------------------------------------------- */

/* a4+a5+a23  BaseKernel.__call__                                   ppi/policies.py:735-745
 * x[b] = mean_fn + mean + l_in z[b] l_out^T, clipped.  z,x: (K,H,d); l_in (H,H); l_out (d,d).
 * l_out_is_diag != 0 promises l_out is diagonal (design_moments, ppi/policies.py:40-50).    */
int ppi_sample_matnorm(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                       int64_t K, int H, int d, const float* mean_fn, const float* mean,
                       const float* l_in, const float* l_out, int l_out_is_diag,
                       const float* lower, const float* upper, float* x, void* stream);

/* a6  BaseFeatures.__call__                                        ppi/policies.py:404-421
 * w[b] = mean + l_in z[b] l_out^T  (K,F,d);  y[b] = [mean_fn + feat w[b], dfeat w[b]] clipped,
 * (K,T,d) or (K,T,2d) when dfeat != NULL.  feat, dfeat: (T,F).  lower/upper have dy entries. */
int ppi_sample_features(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                        int64_t K, int F, int d, int T, const float* mean_fn, const float* mean,
                        const float* l_in, const float* l_out, const float* feat,
                        const float* dfeat, const float* lower, const float* upper, float* w,
                        float* y, void* stream);

/* a8  WhiteNoiseIid.__call__                                       ppi/policies.py:1166-1174 */
int ppi_sample_iid(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset, int64_t K,
                   int H, int d, const float* mean_fn, const float* mean, const float* std,
                   const float* lower, const float* upper, float* x, void* stream);

/* a7  GaussianPolicy.__call__ in Cholesky form  x = mu + z L^T      ppi/policies.py:247-249,
 * ppi/samplers.py:38-39.  L (D,D) lower triangular.  D <= 128 runs an fp32 SIMT kernel; larger
 * D runs the persistent tcgen05 split-product GEMM (PPI_OPT_GEMM_SCHEME) and needs the workspace. */
size_t ppi_sample_gaussian_workspace_bytes(int64_t K, int64_t D);
int ppi_sample_gaussian(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                        int64_t K, int64_t D, const float* mu, const float* L, float* x,
                        void* workspace, size_t workspace_bytes, void* stream);

/* a7 + a16  the same draw, keeping the base normals for the whitened moment match below.
 * zstore (caller-owned, 1 KB aligned, ppi_zstore_bytes(K, D, z != NULL) bytes) receives a 1 KB header
 * (operand scales of this call) and the K rows of base normals as tensor-core operands: one exact
 * fp16 per value for Philox normals, an fp16 pair (scaled by a power of two) for injected ones.
 * D > 128 and PPI_GEMM_FP16X2 only.  x is written exactly as by ppi_sample_gaussian.         */
size_t ppi_zstore_bytes(int64_t K, int64_t D, int injected);
int ppi_sample_gaussian_keep(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset,
                             int64_t K, int64_t D, const float* mu, const float* L, float* x,
                             void* zstore, size_t zstore_bytes, void* workspace,
                             size_t workspace_bytes, void* stream);

/* a7 + a9 fused: ppi_sample_gaussian_keep with a separable test function (PPI_FN_RASTRIGIN / PPI_FN_STYBLINSKI,
 * optimization/functions.py:65-66,78-83) evaluated in the sampling epilogue -- the step `costs = f(actions)`
 * of ppi/algorithms.py:85-86 without a second pass over the K x D samples.  cost (K) fp32.  x may be NULL: the
 * samples are then never written (the whitened update reads the base-normal store; x is re-drawn from the same
 * (seed, draw_id, row_offset) with ppi_sample_gaussian when a caller asks for it).
 * store_ready != 0 (Philox only): zstore already holds this draw (ppi_philox_prefetch), only the GEMMs run.
 * workspace: ppi_sample_gaussian_cost_workspace_bytes(K, D). */
size_t ppi_sample_gaussian_cost_workspace_bytes(int64_t K, int64_t D);
int ppi_sample_gaussian_cost(const float* z, uint64_t seed, uint64_t draw_id, uint64_t row_offset, int64_t K,
                             int64_t D, const float* mu, const float* L, float* x, void* zstore,
                             size_t zstore_bytes, int store_ready, int fn, float* cost, void* workspace,
                             size_t workspace_bytes, void* stream);

/* Base normals of a Philox draw generated ahead of the sampling call (they depend on (seed, draw_id,
 * global row) only): zstore = the K-major store of ppi_sample_gaussian_keep, 1 KB aligned, size from
 * ppi_zstore_bytes(K, D, 0).  Issued on a side stream this hides the generator (ppi/samplers.py:18-20,
 * np.random.randn) under the HBM- and latency-bound part of the previous update.
 * ppi_sample_gaussian_pre is ppi_sample_gaussian_keep for such a draw: only the GEMM passes run. */
int ppi_philox_prefetch(uint64_t seed, uint64_t draw_id, uint64_t row_offset, int64_t K, int64_t D,
                        void* zstore, size_t zstore_bytes, void* stream);
int ppi_sample_gaussian_pre(int64_t K, int64_t D, const float* mu, const float* L, float* x, void* zstore,
                            void* workspace, size_t workspace_bytes, void* stream);

/* The base normals the dense Gaussian sampler (D > 128: ppi_sample_gaussian / _keep / _cost / _pre) draws for
 * (seed, draw_id, row_offset), as a row-major fp32 matrix (rows, cols).  That sampler keeps its normals K-major
 * (one store serves the sampling product and the whitened covariance), so its Philox counters run down the
 * COLUMNS: one call = samples 4Q..4Q+3 of one column (counter (Q, draw_id << 32 | column), key seed ^ tag).
 * ppi_philox_normal is the row-quad order of every other sampler.  Same fp16-grid values, same guarantees:
 * element (r, c) is a pure function of (seed, draw_id, row_offset + r, c). */
int ppi_philox_normal_cq(float* out, int64_t rows, int64_t cols, uint64_t seed, uint64_t draw_id,
                         uint64_t row_offset, void* stream);

/* ------------------------------------------------------------------------------------------
 * a9  cost evaluation                                     optimization/functions.py:21-104
 * fn: 0 Himmelblau, 1 Rosenbrock, 2 Styblinski, 3 Rastrigin, 4 NoisySphere (aux = A (D,D),
 * noise = (K,) standard normals or NULL).  x (K,D) -> c (K,).
 * ---------------------------------------------------------------------------------------- */
enum { PPI_FN_HIMMELBLAU = 0, PPI_FN_ROSENBROCK = 1, PPI_FN_STYBLINSKI = 2, PPI_FN_RASTRIGIN = 3,
       PPI_FN_NOISY_SPHERE = 4 };
int ppi_cost_function(int fn, const float* x, int64_t K, int64_t D, const float* aux,
                      const float* noise, float* c, void* stream);

/* synthetic trajectory return = the Test env tracking cost     policy_search/envs.py:65-72,83-87
 * c[b] = sum_{t, j<dt} (x[b,t,j] - target[t,j])^2 / (T*dt);  x (K,T,dy), dy >= dt.          */
int ppi_cost_tracking(const float* x, int64_t K, int T, int dy, const float* target, int dt,
                      float* c, void* stream);

/* ------------------------------------------------------------------------------------------
 * Weights.  All results of one update that the host reads come back in one struct.
 * ---------------------------------------------------------------------------------------- */
typedef struct ppi_weight_stats {
  double alpha;          /* inverse temperature used                                         */
  double lse;            /* log sum_k exp(log_w_k)                                           */
  double ess;            /* exp(-LSE(2 log nw))                 ppi/policies.py:87           */
  double weighted_cost;  /* sum_k nw_k c_k on the normalised costs                           */
  double weight_entropy; /* sum_k nw_k log nw_k                 ppi/algorithms.py:123,622    */
  double cost_min, cost_max, cost_mean, cost_std; /* over non-NaN costs (:89-90, :184)       */
  double objective;      /* value of the solved 1-D objective at alpha (dual / bound / |ess-N|) */
  int64_t n_valid;       /* non-NaN costs; == K when all are NaN (all-NaN keeps everything)  */
  int64_t argmax;        /* first index of the largest weight   ppi/policies.py:811          */
  int32_t all_nan;       /* 1 when every cost was NaN           ppi/algorithms.py:62-64      */
  int32_t iterations;    /* objective evaluations spent by the solver                        */
  int32_t status;        /* 0 ok, 1 bracket hit a bound, 2 max iterations                    */
  int32_t pad;
} ppi_weight_stats;

/* a2+a10..a14  NaN mask, normalisation, temperature solve, normalised weights.
 * ppi/algorithms.py:57-68 (filter), :184,:510,:541,:597 (normalise), :472 (MPPI shift),
 * :183-215 (REPS dual), :539-567 (LBPS), :596-615 (ESSPS), :509-511 (AIS).
 * mode / hyper:
 *   PPI_TEMP_FIXED_NORMALISED  hyper = alpha     log_w = -alpha (c-min)/(max-min)      (Ais)
 *   PPI_TEMP_FIXED_SHIFTED     hyper = alpha     log_w = -alpha (c-min)                (Mppi)
 *   PPI_TEMP_REPS_KL           hyper = epsilon   KL(nw || uniform) = epsilon, alpha in [1e-5,5e2]
 *   PPI_TEMP_ESS_TARGET        hyper = n_elites  ESS(alpha) = n_elites
 *   PPI_TEMP_LOWER_BOUND       hyper = delta     argmin_alpha E_nw[c] + sqrt((1-delta)/delta)/sqrt(ESS)
 * NaN costs get weight 0 (the reference compacts them away); if all are NaN every cost is 0.
 * nw (K) receives normalised weights, log_w (K, may be NULL) the unnormalised log-weights with
 * -inf at NaN costs.  workspace: ppi_solve_temperature_workspace_bytes(K).                  */
enum { PPI_TEMP_FIXED_NORMALISED = 0, PPI_TEMP_FIXED_SHIFTED = 1, PPI_TEMP_REPS_KL = 2,
       PPI_TEMP_ESS_TARGET = 3, PPI_TEMP_LOWER_BOUND = 4 };
size_t ppi_solve_temperature_workspace_bytes(int64_t K);
int ppi_solve_temperature(const float* costs, int64_t K, int mode, double hyper, float* nw,
                          float* log_w, ppi_weight_stats* stats, void* workspace,
                          size_t workspace_bytes, void* stream);
/* ppi_solve_temperature without a workspace argument and without any allocation.  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, results and error
 * behaviour as ppi_solve_temperature; the solver's scratch (grid-barrier counter + per-block partial sums,
 * 75 904 bytes) is a file-scope `static __device__` array inside the library, held by every GPU that loads it
 * until the process exits.  Limits and rules:
 *   - K <= PPI_SOLVE_STATIC_MAX_K costs, larger K is refused with PPI_ERR_UNSUPPORTED;
 *   - calls no cudaMalloc / cudaFree / allocating routine and creates no CUDA events (it is not bracketed by
 *     ppi_profile_enable); the FIRST call on each device resolves the array's device address
 *     (cudaGetSymbolAddress, once per device, with the caller's current device) -- make one call per device
 *     before capturing a graph; `stream` must belong to the current device;
 *   - every call on a device shares the one array: calls that overlap on the GPU (different streams, or graph
 *     branches without a dependency) race -- serialise them, or use ppi_solve_temperature with one workspace
 *     per stream.  The array keeps the previous call's contents; what must start at zero (the barrier counter)
 *     is cleared on `stream` at the start of every call. */
#define PPI_SOLVE_STATIC_MAX_K (1 << 22)
int ppi_solve_temperature_static(const float* costs, int64_t K, int mode, double hyper, float* nw,
                                 float* log_w, ppi_weight_stats* stats, void* stream);

/* user-supplied log-weights -> normalised weights (m_projection head)  ppi/policies.py:84-87,
 * argmax ppi/policies.py:811,1153.  costs may be NULL (then the cost_* fields are 0).       */
int ppi_normalise_log_weights(const float* log_w, int64_t K, float* nw, ppi_weight_stats* stats,
                              void* workspace, size_t workspace_bytes, void* stream);

/* a15  CEM / iCEM elites                                         ppi/algorithms.py:117-120,149-154
 * sorted_idx (K) = argsort by (cost, index), NaN last; nw = 1/k on the first k, 0 elsewhere
 * (log_w in {0,-1e12} => exactly 1/k).  workspace: ppi_cem_workspace_bytes(K).              */
size_t ppi_cem_workspace_bytes(int64_t K);
int ppi_cem_weights(const float* costs, int64_t K, int64_t n_elites, int64_t* sorted_idx,
                    float* nw, ppi_weight_stats* stats, void* workspace, size_t workspace_bytes,
                    void* stream);

/* ------------------------------------------------------------------------------------------
 * Moment matching.
 * ---------------------------------------------------------------------------------------- */

/* a16/a18/a21 first moment and element-wise second moment, HBM-bound float4 reduction.
 * s1[i] += sum_k nw_k (x_ki - shift1_i);   s2[i] += sum_k nw_k (x_ki - shift2_i)^2  (s2 may be
 * NULL; shifts may be NULL = 0).  x is (K,D) with D = H*d for trajectories.  fp64 accumulators
 * must be zeroed by the caller (K-sharded ranks all-reduce them).  absdev_max (device scalar,
 * may be NULL, zeroed by the caller) receives max |x_ki - shift1_i| over the rows with nw_k != 0
 * (max-combined, so K-sharded ranks all-reduce it with MAX): twice this value bounds the
 * deviation from the NEW mean and is what ppi_wmoments_dense takes as dev_bound.
 * ppi/policies.py:88 (mu), :122 (matrix-normal mean), :1158-1162 (iid mean / std).          */
int ppi_wmoments_diag(const float* x, const float* nw, int64_t K, int64_t D, const float* shift1,
                      const float* shift2, double* s1, double* s2, float* absdev_max, void* stream);

/* a18  matrix-normal input covariance                               ppi/policies.py:129-132
 * acc[i,l] += sum_b nw_b sum_j (x_bij - shift_ij)(x_blj - shift_lj) * inv_cov_out_j
 * (caller divides by d).  x (K,H,d); shift (H,d); inv_cov_out (d) or NULL (= 1).  H <= 128.
 * With d = 1 this is the small-D dense covariance of ppi/policies.py:91.                   */
int ppi_wmoments_matnorm(const float* x, const float* nw, int64_t K, int H, int d,
                         const float* shift, const float* inv_cov_out, double* acc, void* stream);

/* a16  dense weighted covariance  sum_k nw_k (x_k - shift)(x_k - shift)^T   ppi/policies.py:89-92
 * on tcgen05 tensor cores (split products per PPI_OPT_GEMM_SCHEME, chunked fp32 promotion),
 * operands staged K-major by a transposing pre-pass and fed by TMA.  cov (D,D) receives the
 * symmetrised result (+= when accumulate != 0).  Any D >= 1; D < 64 is cheaper through
 * ppi_wmoments_matnorm.  dev_bound (device scalar, may be NULL): an upper bound of |x_ki - shift_i|
 * over the rows with nw_k != 0, used by the FP16X2 scheme to pick its operand scale; any bound
 * within ~2^10 of the true maximum is as good as the maximum itself.  NULL = measured with one
 * extra pass over x.  ppi_wmoments_diag can produce it for free (absdev_max).               */
size_t ppi_wmoments_dense_workspace_bytes(int64_t K, int64_t D);
int ppi_wmoments_dense(const float* x, const float* nw, int64_t K, int64_t D, const float* shift,
                       const float* dev_bound, float* cov, int accumulate, void* workspace,
                       size_t workspace_bytes, void* stream);

/* a16 in whitened coordinates                                       ppi/policies.py:84-92
 * For the samples x_k = mu + L z_k of ppi_sample_gaussian_keep the M-projection factorises:
 *   sum_k nw_k x_k = mu + L m,                       m = sum_k nw_k z_k
 *   sum_k nw_k (x_k - mu_new)(x_k - mu_new)^T = L (S - m m^T) L^T,   S = sum_k nw_k z_k z_k^T
 * so the K-long contraction runs on the stored base normals: operand A = nw z (fp16 pair), operand
 * B = z -- one exact fp16 for Philox normals, i.e. two partial products instead of three and no
 * pass over x.  s1 (D, fp64, zeroed by the caller) += m;  cov (D,D) receives S symmetrised (+= when
 * accumulate != 0).  s2 (D, fp64, zeroed by the caller; may be NULL) += diag(S) = sum_k nw_k z_k^2 with
 * round-to-nearest products and fp64 sums: tcgen05's fp32 accumulation truncates, which biases sums of positive
 * terms by ~2e-6, and tr C enters the KL as (tr C - D); ppi_white_finish takes the diagonal of C from s2.
 * K-sharded ranks all-reduce all three.  `injected` must match the store.
 * The caller finishes with D x D algebra: chol(new sigma) = L chol(S - m m^T),
 * KL(new || old) = (tr(C) + |m|^2 - D - log det C) / 2 with C = S - m m^T.                   */
size_t ppi_wmoments_white_workspace_bytes(int64_t K, int64_t D, int injected);
int ppi_wmoments_white(const void* zstore, int injected, const float* nw, int64_t K, int64_t D,
                       double* s1, double* s2, float* cov, int accumulate, void* workspace,
                       size_t workspace_bytes, void* stream);

/* a17 tail: sigma = cov - delta delta^T (delta = new mean - shift), optional factorized() and
 * symmetric()                                                        ppi/policies.py:53-62,92 */
int ppi_finalize_gaussian(const double* s1, const float* cov_in, int64_t D, const float* shift,
                          int diagonal_only, float* mu_out, float* sigma_out, void* stream);

/* a19/a20 tail: mean = s1 (H,d); cov_in = sym(acc)/d - correction; fp64 Cholesky (one block).
 * jitter is added to the diagonal before factorising (1e-12 for features, policies.py:480). */
int ppi_finalize_matnorm(const double* s1, const double* acc, int H, int d, double jitter,
                         double* mean64, float* mean32, double* cov64, float* cov32,
                         double* chol64, float* chol32, int* status, void* stream);

/* ------------------------------------------------------------------------------------------
 * f3  dense D x D fp64 tail of the Gaussian moment match, any D (blocked, fp64 tensor cores)
 *     np.linalg.cholesky + PD check                  ppi/policies.py:259-268; ppi/samplers.py:38-39
 *     log det / trace terms of KL and entropy        ppi/policies.py:146-158,213-216
 * scalars (4 doubles, device): [0] status (0 = positive definite, else 1 + index of the first failing
 * pivot -- np.linalg.LinAlgError), [1] log det of the factored matrix, [2] tr C, [3] |m|^2 (white finish).
 * workspace: ppi_dense_workspace_bytes(D) (three padded D x D fp64 matrices).
 *
 * ppi_cholesky_blocked: chol64 / chol32 (D, D) = lower Cholesky factor of the symmetric a (D, D) fp64 (lower
 *   triangle read), zero above the diagonal; either output may be NULL.
 * ppi_white_finish: the whole tail of the whitened update (ppi_wmoments_white) in one call:
 *   C = S - m m^T (its diagonal from `diag` = s2 of ppi_wmoments_white when non-NULL),  G = chol(C),
 *   l_new = l_old G (fp64 + fp32 copy for the sampler),  mu_new = mu_old + l_old m,
 *   log det C, tr C, |m|^2  ->  KL(new || old) = (tr C + |m|^2 - D - log det C) / 2 on the host.
 *   Tiles above the block diagonal of l_new64 / l_new32 are not written: pass zeroed buffers.
 *   On a failed factorisation l_new / mu_new are left untouched (the caller takes the reference's
 *   regularisation branch).
 * ppi_tri_gram: out = l l^T for a lower-triangular l (sigma from its factor; triangular tile skipping).
 * ---------------------------------------------------------------------------------------- */
size_t ppi_dense_workspace_bytes(int64_t D);
int ppi_cholesky_blocked(const double* a, int64_t D, double* chol64, float* chol32, double* scalars,
                         void* workspace, size_t workspace_bytes, void* stream);
int ppi_white_finish(const float* S, const double* m, const double* diag, const double* l_old, const float* mu_old,
                     int64_t D, double* l_new64, float* l_new32, float* mu_new32, double* scalars, void* workspace,
                     size_t workspace_bytes, void* stream);
int ppi_tri_gram(const double* l, int64_t D, double* out64, float* out32, void* stream);
/* y32 = x0 + l v for a lower-triangular fp64 l (D, D); x0 may be NULL (the mean step mu + L m alone) */
int ppi_tri_matvec(const double* l, int64_t D, const double* v, const float* x0, float* y32, void* stream);
/* The two solve-dependent terms of the Gaussian KL (ppi/policies.py:146-158: np.linalg.solve(sigma_2, .)) from
 * the Cholesky factors, by blocked forward substitution on the fp64 tensor cores:
 *   out[0] = |l^-1 b_lower|_F^2   (= tr(Sigma_old^-1 Sigma_new) for l = chol(Sigma_old), b_lower = chol(Sigma_new))
 *   out[1] = |l^-1 v|^2           (= (mu_1 - mu_2)^T Sigma_old^-1 (mu_1 - mu_2))
 * l, b_lower: (D, D) fp64 lower triangular with zeros above the diagonal; v: (D) fp64; b_lower or v may be NULL.
 * out: 2 doubles on the device.  workspace: ppi_dense_workspace_bytes(D). */
int ppi_tri_solve_sumsq(const double* l, int64_t D, const double* b_lower, const double* v, double* out,
                        void* workspace, size_t workspace_bytes, void* stream);

/* a18-a20 fused for one rank: ppi_wmoments_diag -> centre -> ppi_wmoments_matnorm -> ppi_finalize_matnorm
 * back to back, the result packed for a single device->host read (ppi/policies.py:96-143,804-848,461-516):
 *   packed (fp64) = [ mean (H d) = sum nw (x - shift1) | cov_in (H H) | chol (H H) | status |
 *                     x[map_index] (H d, only when map_index >= 0 or stats: the argmax sample, :811) |
 *                     *stats as raw 8-byte words (only when stats != NULL) ]
 * stats: the ppi_weight_stats a preceding ppi_solve_temperature left ON THE DEVICE; the sample row is then
 * x[stats->argmax] and the statistics travel home in the same read (no synchronisation between the
 * temperature solve and the match).
 * shift1 (H, d) may be NULL (feature weights are matched without a mean function); inv_cov_out (d).
 * mean32 (H, d) / chol32 (H, H), optional: fp32 device copies of the fitted mean and factor -- what the next
 * ppi_sample_matnorm / ppi_sample_features call consumes, so the policy state never has to be uploaded again.
 * workspace: ppi_matnorm_update_workspace_bytes(H, d).  K-sharded ranks use the three separate entry
 * points with their all-reduces in between. */
size_t ppi_matnorm_update_workspace_bytes(int H, int d);
int ppi_matnorm_update(const float* x, const float* nw, int64_t K, int H, int d, const float* shift1,
                       const float* inv_cov_out, double jitter, int64_t map_index,
                       const ppi_weight_stats* stats, double* packed, float* mean32, float* chol32,
                       void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Measurement hooks (new; the reference has no instrumentation beyond profile_mpc.py:96-108).
 * ppi_launch_count: kernels this library has launched since load.  ppi_profile_enable(1) makes
 * the library bracket its kernels with CUDA events on the launching stream; ppi_profile_read
 * returns the number of bracketed scopes of a slot and their summed device time, then
 * ppi_profile_enable(0/1) clears the records.
 * ---------------------------------------------------------------------------------------- */
enum { PPI_PROF_SYRK_GEMM = 0, PPI_PROF_SAMPLE_GEMM = 1, PPI_PROF_SYRK_OPERAND = 2,
       PPI_PROF_SAMPLE_OPERAND = 3, PPI_PROF_WMOMENTS_DIAG = 4, PPI_PROF_COST = 5,
       PPI_PROF_SOLVE = 6, PPI_PROF_SYRK_REDUCE = 7, PPI_PROF_SAMPLE_SMALL = 8,
       PPI_PROF_WMOMENTS_MATNORM = 9, PPI_PROF_DENSE_TAIL = 10, PPI_PROF_NUM_SLOTS = 11 };
int64_t ppi_launch_count(void);
int ppi_profile_enable(int on);
int ppi_profile_read(int slot, int64_t* count, double* total_ms);

#ifdef __cplusplus
}
#endif
#endif /* PPI_B200_H */
