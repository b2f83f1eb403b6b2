// Library-level entry points: version, error string, device check.
#include <atomic>
#include <cstdlib>
#include <stdarg.h>
#include <string.h>
#include "common.cuh"

namespace tdssf {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int current_device() {
  int dev = 0;
  return cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < kMaxDevices ? dev : 0;
}

// per device: one process may drive several GPUs (one library instance, kernels launched on the current device)
int num_sms() {
  static std::atomic<int> cached[kMaxDevices];
  const int dev = current_device();
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
#ifdef TDSSF_EXPERIMENT
    // experiment builds only: TDSSF_SMS=<n> sizes every persistent grid (and the workspaces derived from it) for fewer
    // SMs, to tell a per-SM limit from a chip-wide one (profiles/r1_experiments_v10.md)
    if (const char* e = getenv("TDSSF_SMS")) {
      const int v = atoi(e);
      if (v >= 2 && v < n) n = v & ~1;
    }
#endif
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

}  // namespace tdssf

extern "C" int tdssf_version(void) { return TDSSF_VERSION; }

extern "C" const char* tdssf_last_error(void) { return tdssf::g_err; }

extern "C" int tdssf_check_device(void) {
  int dev = 0, major = 0;
  TDSSF_CUDA(cudaGetDevice(&dev));
  TDSSF_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) {
    tdssf::set_error("tdssf is built for sm_100a only; current device is sm_%d*", major * 10);
    return TDSSF_EDEVICE;
  }
  return TDSSF_OK;
}
