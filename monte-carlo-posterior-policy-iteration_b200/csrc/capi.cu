// Error plumbing + version of the C ABI (include/ppi_b200.h).
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace ppi {
static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}
}  // namespace ppi

namespace ppi {
// ---- launch counter + optional per-kernel CUDA-event timing (bench.py's roofline leg) ----
static std::atomic<int64_t> g_launches{0};
static bool g_profile = false;
struct EventPair { cudaEvent_t a, b; };
static std::vector<EventPair> g_events[PPI_PROF_NUM_SLOTS];
static std::mutex g_prof_mutex;

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

ProfileScope::ProfileScope(int slot, cudaStream_t stream) : slot_(slot), stream_(stream), on_(g_profile) {
  if (!on_) return;
  EventPair ev;
  if (cudaEventCreate(&ev.a) != cudaSuccess || cudaEventCreate(&ev.b) != cudaSuccess) { on_ = false; return; }
  cudaEventRecord(ev.a, stream_);
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  g_events[slot_].push_back(ev);
  index_ = g_events[slot_].size() - 1;
}
ProfileScope::~ProfileScope() {
  if (!on_) return;
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  cudaEventRecord(g_events[slot_][index_].b, stream_);
}
}  // namespace ppi

namespace ppi {
static std::atomic<int> g_gemm_scheme{PPI_GEMM_FP16X2};
int gemm_scheme() { return g_gemm_scheme.load(std::memory_order_relaxed); }
static std::atomic<int> g_gemm_chunk{8};
int gemm_chunk() { return g_gemm_chunk.load(std::memory_order_relaxed); }
static std::atomic<int> g_sample_chunk{16};
int sample_chunk() { return g_sample_chunk.load(std::memory_order_relaxed); }
static std::atomic<int> g_gemm_overlap{0};
int gemm_overlap() { return g_gemm_overlap.load(std::memory_order_relaxed); }
static std::atomic<int> g_gemm_pair{0};
int gemm_pair() { return g_gemm_pair.load(std::memory_order_relaxed); }
static std::atomic<int> g_prefetch_blocks{6};
int prefetch_blocks() { return g_prefetch_blocks.load(std::memory_order_relaxed); }
}  // namespace ppi

extern "C" int ppi_set_option(int key, int value) {
  if (key == PPI_OPT_GEMM_SCHEME) {
    PPI_REQUIRE(value == PPI_GEMM_TF32X3 || value == PPI_GEMM_FP16X2, PPI_ERR_INVALID,
                "ppi_set_option: unknown GEMM scheme %d", value);
    ppi::g_gemm_scheme.store(value);
    return 0;
  }
  if (key == PPI_OPT_GEMM_CHUNK) {
    PPI_REQUIRE(value >= 2 && value <= 4096 && value % 2 == 0, PPI_ERR_INVALID,
                "ppi_set_option: GEMM chunk must be an even number of k-blocks in [2, 4096]");
    ppi::g_gemm_chunk.store(value);
    return 0;
  }
  if (key == PPI_OPT_SAMPLE_CHUNK) {
    PPI_REQUIRE(value >= 2 && value <= 4096 && value % 2 == 0, PPI_ERR_INVALID,
                "ppi_set_option: sampling chunk must be an even number of k-blocks in [2, 4096]");
    ppi::g_sample_chunk.store(value);
    return 0;
  }
  if (key == PPI_OPT_GEMM_OVERLAP) { ppi::g_gemm_overlap.store(value != 0); return 0; }
  if (key == PPI_OPT_GEMM_PAIR) { ppi::g_gemm_pair.store(value != 0); return 0; }
  if (key == PPI_OPT_PREFETCH_BLOCKS) {
    PPI_REQUIRE(value >= 1 && value <= 6, PPI_ERR_INVALID, "ppi_set_option: prefetch CTAs per SM must lie in [1, 6]");
    ppi::g_prefetch_blocks.store(value);
    return 0;
  }
  PPI_REQUIRE(false, PPI_ERR_INVALID, "ppi_set_option: unknown key %d", key);
}

extern "C" int ppi_get_option(int key) {
  return key == PPI_OPT_GEMM_SCHEME ? ppi::gemm_scheme()
       : key == PPI_OPT_GEMM_CHUNK ? ppi::gemm_chunk()
       : key == PPI_OPT_GEMM_OVERLAP ? ppi::gemm_overlap()
       : key == PPI_OPT_GEMM_PAIR ? ppi::gemm_pair()
       : key == PPI_OPT_PREFETCH_BLOCKS ? ppi::prefetch_blocks()
       : key == PPI_OPT_SAMPLE_CHUNK ? ppi::sample_chunk() : -1;
}

extern "C" int64_t ppi_launch_count(void) { return ppi::g_launches.load(); }

extern "C" int ppi_profile_enable(int on) {
  std::lock_guard<std::mutex> lock(ppi::g_prof_mutex);
  for (auto& v : ppi::g_events) {
    for (auto& e : v) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    v.clear();
  }
  ppi::g_profile = on != 0;
  return 0;
}

extern "C" int ppi_profile_read(int slot, int64_t* count, double* total_ms) {
  PPI_REQUIRE(slot >= 0 && slot < PPI_PROF_NUM_SLOTS && count && total_ms, PPI_ERR_INVALID,
              "ppi_profile_read: bad arguments");
  std::lock_guard<std::mutex> lock(ppi::g_prof_mutex);
  double ms = 0.0;
  for (auto& e : ppi::g_events[slot]) {
    PPI_CUDA_TRY(cudaEventSynchronize(e.b));
    float t = 0.f;
    PPI_CUDA_TRY(cudaEventElapsedTime(&t, e.a, e.b));
    ms += t;
  }
  *count = (int64_t)ppi::g_events[slot].size();
  *total_ms = ms;
  return 0;
}

extern "C" int ppi_version(void) { return PPI_B200_VERSION; }
extern "C" const char* ppi_last_error(void) { return ppi::g_error; }
