// Per-point projection math shared by the stand-alone projector and the fused kernel.
// Restates utils/fusion_model.py:63-84 (projector), :117-127 (rescale), :176-188 (patch index)
// with the reference's CPU fp32 evaluation order:
//   extrinsic row  : ((r0*x + r1*y) + r2*z) + r3        — every mul and add rounded separately
//   intrinsic row  : fma(k2, z, fma(k1, y, k0*x))       — FMA chain
//   divide         : IEEE round-to-nearest
#pragma once
#include "common.cuh"

namespace tdssf {

// Per-sweep calibration staged in shared memory: NC x (12 extrinsic + 9 intrinsic) floats.
struct CamCalib {
  float r[3][4];  // first three rows of lidar->camera
  float k[3][3];
};

struct SweepGeom {
  float img_h, img_w;      // original image size
  float scale_h, scale_w;  // resize / original, fp32 division like the reference
  float patch;             // float(patch_size)
  int grid_w, grid_h, total;
};

__device__ __forceinline__ SweepGeom make_geom(float img_h, float img_w, int resize_h,
                                               int resize_w, int patch) {
  SweepGeom g;
  g.img_h = img_h;
  g.img_w = img_w;
  // scale_pixel_coords: H_new / clamp(H_orig, 1e-6)   (utils/fusion_model.py:117-121)
  g.scale_h = __fdiv_rn((float)resize_h, fmaxf(img_h, 1e-6f));
  g.scale_w = __fdiv_rn((float)resize_w, fmaxf(img_w, 1e-6f));
  g.patch = (float)patch;
  g.grid_h = patch > 0 ? resize_h / patch : 0;
  g.grid_w = patch > 0 ? resize_w / patch : 0;
  g.total = g.grid_h * g.grid_w;
  return g;
}

struct Proj {
  float u, v, z;  // u,v = -1 and z = -1 where invalid (reference outputs)
  bool valid;
};

// Homogeneous image coordinates of one (point, camera): everything of the projector before the perspective divide.
struct Homog {
  float cam_z;       // depth in the camera frame
  float p0, p1;      // K rows 0,1 applied to the camera-frame point
  float den;         // clamp(K row 2 applied, min=1e-12), NaN kept like torch does
};

__device__ __forceinline__ Homog project_homog(const CamCalib& c, float x, float y, float z) {
  float cam[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = __fadd_rn(__fmul_rn(c.r[i][0], x), __fmul_rn(c.r[i][1], y));
    a = __fadd_rn(a, __fmul_rn(c.r[i][2], z));
    cam[i] = __fadd_rn(a, c.r[i][3]);
  }
  float pix[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = __fmul_rn(c.k[i][0], cam[0]);
    a = __fmaf_rn(c.k[i][1], cam[1], a);
    pix[i] = __fmaf_rn(c.k[i][2], cam[2], a);
  }
  Homog h;
  h.cam_z = cam[2];
  h.p0 = pix[0];
  h.p1 = pix[1];
  h.den = (pix[2] != pix[2]) ? pix[2] : fmaxf(pix[2], 1e-12f);
  return h;
}

__device__ __forceinline__ Proj project_finish(const Homog& h, float img_h, float img_w) {
  const float u = __fdiv_rn(h.p0, h.den);
  const float v = __fdiv_rn(h.p1, h.den);
  Proj p;
  p.valid = (h.cam_z > 0.f) && (u >= 0.f) && (u < img_w) && (v >= 0.f) && (v < img_h);
  p.u = p.valid ? u : -1.f;
  p.v = p.valid ? v : -1.f;
  p.z = p.valid ? h.cam_z : -1.f;
  return p;
}

// Division-free filter: false ONLY when project_finish(h).valid is certainly false, so the two IEEE divisions and the
// patch arithmetic run for the ~13 % of (point, camera) pairs that can be visible.  With den >= 1e-12 (or NaN):
//   p0 < -1e-20*den            =>  p0/den <= -0.9e-20, rounds to a negative float              =>  u >= 0 fails
//   p0 > 1.000001*(W*den)      =>  p0/den > W*(1 + 8e-7) after the two roundings, rounds >= W  =>  u <  W fails
// Every comparison is false on NaN, so NaN operands are kept for the exact path, which rejects them.
__device__ __forceinline__ bool maybe_visible(const Homog& h, float img_h, float img_w) {
  const float lo = __fmul_rn(-1e-20f, h.den);
  const float hi_u = __fmul_rn(1.000001f, __fmul_rn(img_w, h.den));
  const float hi_v = __fmul_rn(1.000001f, __fmul_rn(img_h, h.den));
  const bool out = !(h.cam_z > 0.f) || (h.p0 < lo) || (h.p0 > hi_u) || (h.p1 < lo) || (h.p1 > hi_v);
  return !out;
}

__device__ __forceinline__ Proj project_one(const CamCalib& c, float x, float y, float z,
                                            float img_h, float img_w) {
  return project_finish(project_homog(c, x, y, z), img_h, img_w);
}

// (u,v) in original pixels (or -1) -> clamped flat patch index.
__device__ __forceinline__ int patch_index(const SweepGeom& g, float u, float v) {
  float su = __fmul_rn(u, g.scale_w);
  float sv = __fmul_rn(v, g.scale_h);
  // (pixel / float(patch)).long(): IEEE divide, truncate toward zero
  int px = __float2int_rz(__fdiv_rn(su, g.patch));
  int py = __float2int_rz(__fdiv_rn(sv, g.patch));
  px = min(max(px, 0), g.grid_w - 1);
  py = min(max(py, 0), g.grid_h - 1);
  int flat = py * g.grid_w + px;
  return min(max(flat, 0), g.total - 1);
}

// Loads NC cameras of sweep b into shared memory (all threads of the block take part).
__device__ __forceinline__ void load_calib(CamCalib* s_cal, const float* __restrict__ K,
                                           const float* __restrict__ Rt, int b, int NC) {
  for (int i = threadIdx.x; i < NC * 21; i += blockDim.x) {
    int cam = i / 21, j = i % 21;
    float val;
    if (j < 12)
      val = Rt[((int64_t)b * NC + cam) * 16 + j];  // rows 0..2 of the 4x4
    else
      val = K[((int64_t)b * NC + cam) * 9 + (j - 12)];
    reinterpret_cast<float*>(&s_cal[cam])[j] = val;
  }
}

}  // namespace tdssf
