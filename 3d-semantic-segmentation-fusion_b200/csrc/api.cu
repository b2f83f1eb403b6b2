// Library-level entry points: version, error string, device check.
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

int num_sms() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      cached = 148;
    // experiment hook: TDSSF_SMS=<n> sizes every persistent grid (and the workspaces derived from it) for fewer SMs, to
    // tell a per-SM limit from a chip-wide one (profiles/r1_experiments_v10.md)
    if (const char* e = getenv("TDSSF_SMS")) {
      const int v = atoi(e);
      if (v >= 2 && v < cached) cached = v & ~1;
    }
  }
  return cached;
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
