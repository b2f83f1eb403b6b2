// Shared helpers for the tdssf kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/tdssf.h"

namespace tdssf {

void set_error(const char* fmt, ...);

#define TDSSF_REQUIRE(cond, ...)                         \
  do {                                                   \
    if (!(cond)) {                                       \
      ::tdssf::set_error(__VA_ARGS__);                   \
      return TDSSF_EINVAL;                               \
    }                                                    \
  } while (0)

#define TDSSF_CUDA(call)                                                              \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      ::tdssf::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,                 \
                         cudaGetErrorString(e__));                                    \
      return TDSSF_ECUDA;                                                             \
    }                                                                                 \
  } while (0)

#define TDSSF_LAUNCH_CHECK()                                                          \
  do {                                                                                \
    cudaError_t e__ = cudaGetLastError();                                             \
    if (e__ != cudaSuccess) {                                                         \
      ::tdssf::set_error("%s:%d launch -> %s", __FILE__, __LINE__,                    \
                         cudaGetErrorString(e__));                                    \
      return TDSSF_ECUDA;                                                             \
    }                                                                                 \
  } while (0)

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

constexpr int kMaxDevices = 64;
int current_device();  // index of the current CUDA device (0 when it cannot be read), < kMaxDevices
int num_sms();         // SM count of the current device (cached per device)

// streaming (read-once) 16-byte load that does not allocate in L1
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_f4(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

}  // namespace tdssf
