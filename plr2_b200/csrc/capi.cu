// C-ABI housekeeping: version, last-error text, device query.
#include "common.cuh"
#include <stdarg.h>
#include <stdlib.h>

namespace plr2 {
static thread_local char g_err[512] = "";
long long g_launch_count = 0;
void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
bool infer_fp16() {
  static int v = -1;
  if (v < 0) { const char *e = getenv("PLR2_INFER_FMT"); v = (e && e[0] == 'b') ? 0 : 1; }
  return v != 0;
}
int infer_conv6_nprod() {
  static int v = -1;
  if (v < 0) { const char *e = getenv("PLR2_CONV6_NPROD"); v = e ? atoi(e) : 2; if (v != 1) v = 2; }
  return v;
}
int device_sm_count() {
  static PerDeviceInt cache;
  const int dev = current_device();
  int n = cache.get(dev);
  if (!n) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) { cudaGetLastError(); n = 148; }
    cache.set(dev, n);
  }
  return n;
}
}  // namespace plr2

extern "C" int plr2_abi_version(void) { return PLR2_ABI_VERSION; }
extern "C" const char *plr2_last_error_string(void) { return plr2::g_err; }
extern "C" int plr2_device_sm_count(void) {
  int dev = 0, n = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) {
    plr2::set_error("plr2_device_sm_count: %s", cudaGetErrorString(e));
    return -1;
  }
  return n;
}
extern "C" long long plr2_launch_count(void) { return plr2::g_launch_count; }
