// api.cu — error reporting and device queries for libdlion.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace dlion {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
  return (int)e;
}

}  // namespace dlion

extern "C" const char* dlion_last_error(void) { return dlion::g_err; }
extern "C" int dlion_version(void) { return DLION_VERSION; }

extern "C" int dlion_device_sm_count(int* out) {
  int dev = 0, sms = 0;
  int rc = dlion::check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
  if (rc) return rc;
  rc = dlion::check_cuda(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev), "cudaDeviceGetAttribute");
  if (rc) return rc;
  *out = sms;
  return 0;
}
