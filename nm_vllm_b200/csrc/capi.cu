// C-ABI plumbing shared by all kernels: thread-local error string, device queries, build info.
#include "common.cuh"
#include <mutex>
#include <stdlib.h>

namespace b200 {
static thread_local char g_err[512] = "";

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_NO_PDL");
    v = (e && atoi(e) != 0) ? 0 : 1;
  }
  return v == 1;
}

int num_sms() {
  static int cached[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    cached[dev] = v > 0 ? v : 148;
  }
  return cached[dev];
}
}  // namespace b200

extern "C" {

const char* b200_last_error(void) { return b200::g_err; }

const char* b200_build_info(void) {
  return "b200ops sm_100a; cuda " __DATE__;
}

// reference: csrc/cuda_utils_kernels.cu get_device_attribute
int b200_get_device_attribute(int attribute, int device, int* value_out) {
  int dev = device;
  if (device < 0) cudaGetDevice(&dev);
  int v = 0;
  cudaError_t e = cudaDeviceGetAttribute(&v, (cudaDeviceAttr)attribute, dev);
  if (e != cudaSuccess) return b200::set_error(B200_ECUDA, "get_device_attribute: %s", cudaGetErrorString(e));
  *value_out = v;
  return B200_OK;
}

int b200_get_max_shared_memory_per_block_device_attribute(int device, int* value_out) {
  return b200_get_device_attribute((int)cudaDevAttrMaxSharedMemoryPerBlockOptin, device, value_out);
}

}  // extern "C"
