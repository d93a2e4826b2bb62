// Error plumbing + version of the C ABI (include/ppi_b200.h).
#include <stdarg.h>
#include <string.h>

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

extern "C" int ppi_version(void) { return PPI_B200_VERSION; }
extern "C" const char* ppi_last_error(void) { return ppi::g_error; }
