// a1+a2+a3: multi-camera projection, depth/bounds mask, pixel rescale and patch index.
// One thread per LiDAR point, the six cameras unrolled in registers; calibration of the sweep lives
// in shared memory.  HBM-bound: 16 B in, 94 B out (reference layout) or 16 B out (records).
#include <algorithm>
#include "project.cuh"

namespace tdssf {

constexpr int kProjThreads = 256;

template <bool kVec4>
__global__ void __launch_bounds__(kProjThreads)
project_kernel(const float* __restrict__ xyz, int64_t sweep_stride, int64_t point_stride,
               const float* __restrict__ K, const float* __restrict__ Rt,
               const float* __restrict__ img_hw, int V, int NC, int resize_h, int resize_w,
               int patch, float* __restrict__ pix, float* __restrict__ depth,
               uint8_t* __restrict__ valid, int32_t* __restrict__ flat_idx,
               tdssf_rec* __restrict__ rec) {
  __shared__ CamCalib s_cal[16];
  const int b = blockIdx.y;
  load_calib(s_cal, K, Rt, b, NC);
  const SweepGeom g = make_geom(img_hw[2 * b], img_hw[2 * b + 1], resize_h, resize_w, patch);
  __syncthreads();

  const bool want_idx = (flat_idx != nullptr) || (rec != nullptr);
  for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < V; v += gridDim.x * blockDim.x) {
    const float* p = xyz + b * sweep_stride + v * point_stride;
    float x, y, z;
    if (kVec4) {
      float4 q = __ldg(reinterpret_cast<const float4*>(p));
      x = q.x; y = q.y; z = q.z;
    } else {
      x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2);
    }
    uint32_t pk[3] = {0u, 0u, 0u};  // six u16 patch indices
    uint32_t mask = 0, count = 0;
    auto one_cam = [&](int c, uint32_t& patch_out) {
      Proj pr = project_one(s_cal[c], x, y, z, g.img_h, g.img_w);
      const int64_t o = ((int64_t)b * NC + c) * V + v;
      if (pix) reinterpret_cast<float2*>(pix)[o] = make_float2(pr.u, pr.v);
      if (depth) depth[o] = pr.z;
      if (valid) valid[o] = pr.valid ? 1 : 0;
      if (want_idx) {
        int f = patch_index(g, pr.u, pr.v);
        if (flat_idx) flat_idx[o] = f;
        patch_out = (uint32_t)f;
        if (pr.valid) { mask |= 1u << c; count++; }
      }
    };
#pragma unroll
    for (int c = 0; c < TDSSF_MAX_CAMS; ++c) {
      if (c < NC) {
        uint32_t f = 0;
        one_cam(c, f);
        pk[c >> 1] |= (f & 0xffffu) << ((c & 1) * 16);
      }
    }
    for (int c = TDSSF_MAX_CAMS; c < NC; ++c) {  // reference-layout outputs only
      uint32_t f = 0;
      one_cam(c, f);
    }
    if (rec)
      reinterpret_cast<uint4*>(rec)[(int64_t)b * V + v] =
          make_uint4(pk[0], pk[1], pk[2], (mask & 0xffu) | ((count & 0xffu) << 8));
  }
}

// a2 stand-alone: out = pix * (W_new/W_orig, H_new/H_orig), one float2 per thread.
__global__ void scale_pixels_kernel(const float2* __restrict__ pix, const float* __restrict__ img_hw, int hw_rows,
                                    int64_t per_sweep, int64_t total, float new_h, float new_w,
                                    float2* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = hw_rows == 1 ? 0 : (int)(i / per_sweep);
    const float sh = __fdiv_rn(new_h, fmaxf(__ldg(img_hw + 2 * b), 1e-6f));
    const float sw = __fdiv_rn(new_w, fmaxf(__ldg(img_hw + 2 * b + 1), 1e-6f));
    const float2 q = pix[i];
    out[i] = make_float2(__fmul_rn(q.x, sw), __fmul_rn(q.y, sh));
  }
}

}  // namespace tdssf

extern "C" int tdssf_scale_pixels(const float* pix, const float* img_hw, int hw_rows, int B, int64_t per_sweep,
                                  float new_h, float new_w, float* out, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(pix && img_hw && out, "tdssf_scale_pixels: NULL pointer");
  TDSSF_REQUIRE(hw_rows == 1 || hw_rows == B, "tdssf_scale_pixels: origin sizes have %d rows for a batch of %d", hw_rows, B);
  TDSSF_REQUIRE((((uintptr_t)pix | (uintptr_t)out) & 7) == 0, "tdssf_scale_pixels: 8-byte alignment required");
  const int64_t total = (int64_t)B * per_sweep;
  if (total <= 0) return TDSSF_OK;
  const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div(total, 256), (int64_t)num_sms() * 16);
  scale_pixels_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float2*>(pix), img_hw, hw_rows,
                                                                per_sweep, total, new_h, new_w,
                                                                reinterpret_cast<float2*>(out));
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

extern "C" int tdssf_project(const float* xyz, int64_t sweep_stride, int64_t point_stride,
                             const float* K, const float* Rt, const float* img_hw, int B, int V,
                             int NC, int resize_h, int resize_w, int patch, float* pix,
                             float* depth, uint8_t* valid, int32_t* flat_idx, tdssf_rec* rec,
                             void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(xyz && K && Rt && img_hw, "tdssf_project: NULL input");
  TDSSF_REQUIRE(B >= 0 && V >= 0, "tdssf_project: negative size");
  TDSSF_REQUIRE(NC >= 1 && NC <= 16, "tdssf_project: NC=%d not in 1..16", NC);
  TDSSF_REQUIRE(point_stride >= 3, "tdssf_project: point_stride %lld < 3", (long long)point_stride);
  const bool want_idx = flat_idx || rec;
  if (want_idx) {
    TDSSF_REQUIRE(patch > 0 && resize_h >= patch && resize_w >= patch,
                  "tdssf_project: bad resize (%d,%d) / patch %d", resize_h, resize_w, patch);
    TDSSF_REQUIRE((int64_t)(resize_h / patch) * (resize_w / patch) <= 65536,
                  "tdssf_project: patch grid larger than 65536");
  }
  if (rec) {
    TDSSF_REQUIRE(NC <= TDSSF_MAX_CAMS, "tdssf_project: records hold at most %d cameras", TDSSF_MAX_CAMS);
    TDSSF_REQUIRE(((uintptr_t)rec & 15) == 0, "tdssf_project: rec not 16-byte aligned");
  }
  if (pix) TDSSF_REQUIRE(((uintptr_t)pix & 7) == 0, "tdssf_project: pix not 8-byte aligned");
  if (B == 0 || V == 0) return TDSSF_OK;
  const bool vec4 = (point_stride == 4) && (sweep_stride % 4 == 0) && (((uintptr_t)xyz & 15) == 0);
  dim3 grid((unsigned)ceil_div(V, kProjThreads), (unsigned)B);
  cudaStream_t s = (cudaStream_t)stream;
  if (vec4)
    project_kernel<true><<<grid, kProjThreads, 0, s>>>(xyz, sweep_stride, point_stride, K, Rt, img_hw,
                                                       V, NC, resize_h, resize_w, patch, pix, depth,
                                                       valid, flat_idx, rec);
  else
    project_kernel<false><<<grid, kProjThreads, 0, s>>>(xyz, sweep_stride, point_stride, K, Rt,
                                                        img_hw, V, NC, resize_h, resize_w, patch, pix,
                                                        depth, valid, flat_idx, rec);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}
