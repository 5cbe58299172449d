// Error plumbing and version of the C ABI (include/zonos_b200.h).
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/zonos_b200.h"
#include "common.cuh"

namespace zb {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return ZB200_ERR_CUDA;
}
bool pdl_enabled() {
  static const bool on = (zb::tuning_env("ZB200_NO_PDL") == nullptr);
  return on;
}
}  // namespace zb

extern "C" int zb200_version(void) { return ZB200_ABI_VERSION; }
extern "C" const char* zb200_last_error(void) { return zb::g_err; }
