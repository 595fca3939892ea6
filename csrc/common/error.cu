// Last-error slot of the C ABI (thread-local) plus version / capability queries.
#include "util.cuh"
#include <stdarg.h>

namespace hyp {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }
}  // namespace hyp

HYP_API const char* hyp_last_error() { return hyp::get_error(); }
HYP_API int hyp_version() { return 100; }
// 1 when the current device is compute capability 10.x (sm_100a cubins load).
HYP_API int hyp_device_supported() {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}
