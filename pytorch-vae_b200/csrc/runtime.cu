// Error plumbing, version and device-property cache of libvaekit.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace vk {

char* err_buf() {
  static thread_local char buf[512] = "";
  return buf;
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

int check_launch(const char* what) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();  // clear the (non-sticky) launch error
    return fail(VK_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  }
  return VK_OK;
}

int sm_count() {
  // immutable per-device property cache (the only global state of the library)
  static thread_local int cached_dev = -1, cached_sms = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached_dev = dev;
    cached_sms = n;
  }
  return cached_sms;
}

}  // namespace vk

extern "C" int vk_version(void) { return 100; }
extern "C" const char* vk_last_error_string(void) { return vk::err_buf(); }
extern "C" int64_t vk_workspace_bytes(void) { return vk::kWsBytes; }
