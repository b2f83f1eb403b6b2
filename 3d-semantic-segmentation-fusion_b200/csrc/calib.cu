// f4: calibration composition and multi-sweep accumulation on the device (input preparation upstream of a1).
//
// tdssf_compose_lidar2cam replaces the numpy block of utils/dataloader.py:123-152: per camera
//   lidar2cam = inv(T_cam->ego) @ T_lidar->ego            (float64; cast to fp32 at utils/dataloader.py:226)
// from the nuScenes calibrated_sensor records (quaternion w,x,y,z + translation).  The reference also fetches the two
// ego_pose records and never uses them; with them the chain becomes north_star's four stages
//   lidar -> ego(t_lidar) -> global -> ego(t_cam) -> camera.
// tdssf_compose_sweep2ref / tdssf_accumulate_sweeps build the 10-sweep cloud of BASELINE configs[3] the way
// nuscenes-devkit's LidarPointCloud.from_file_multisweep does (ref_from_car . car_from_global . global_from_car .
// car_from_current, float64, result stored as fp32); the reference itself only ever loads one sweep.
// Everything is a few hundred flops per pose and 32 bytes per point: launch-latency / HBM bound, no tensor-core stage.
#include "common.cuh"

namespace tdssf {
namespace calib {

struct M4 { double m[12]; };  // rows 0..2 of a rigid 4x4 (row 3 = 0 0 0 1)

// pose record: q = (w,x,y,z), t = (x,y,z).  Rotation as pyquaternion.Quaternion.rotation_matrix: the unit quaternion's
// Q . Qbar^T product, whose diagonal is w^2 + x^2 - y^2 - z^2 etc.
__device__ __forceinline__ M4 from_pose(const double* p) {
  double w = p[0], x = p[1], y = p[2], z = p[3];
  const double n = sqrt(w * w + x * x + y * y + z * z);
  if (n > 0.0) { w /= n; x /= n; y /= n; z /= n; }
  M4 r;
  r.m[0] = w * w + x * x - y * y - z * z; r.m[1] = 2.0 * (x * y - w * z);         r.m[2] = 2.0 * (x * z + w * y);         r.m[3] = p[4];
  r.m[4] = 2.0 * (x * y + w * z);         r.m[5] = w * w - x * x + y * y - z * z; r.m[6] = 2.0 * (y * z - w * x);         r.m[7] = p[5];
  r.m[8] = 2.0 * (x * z - w * y);         r.m[9] = 2.0 * (y * z + w * x);         r.m[10] = w * w - x * x - y * y + z * z; r.m[11] = p[6];
  return r;
}
// inverse of a rigid transform: [R^T | -R^T t]
__device__ __forceinline__ M4 rigid_inverse(const M4& a) {
  M4 r;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) r.m[4 * i + j] = a.m[4 * j + i];
    r.m[4 * i + 3] = -(a.m[i] * a.m[3] + a.m[4 + i] * a.m[7] + a.m[8 + i] * a.m[11]);
  }
  return r;
}
__device__ __forceinline__ M4 mul(const M4& a, const M4& b) {
  M4 r;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      double s = a.m[4 * i] * b.m[j] + a.m[4 * i + 1] * b.m[4 + j] + a.m[4 * i + 2] * b.m[8 + j];
      if (j == 3) s += a.m[4 * i + 3];
      r.m[4 * i + j] = s;
    }
  }
  return r;
}
__device__ __forceinline__ void store16(const M4& t, float* o32, double* o64) {
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const double v = k < 12 ? t.m[k] : (k == 15 ? 1.0 : 0.0);
    if (o32) o32[k] = (float)v;
    if (o64) o64[k] = v;
  }
}

__global__ void compose_lidar2cam_kernel(const double* __restrict__ lidar_sensor, const double* __restrict__ cam_sensor,
                                         const double* __restrict__ lidar_ego, const double* __restrict__ cam_ego,
                                         int B, int NC, float* __restrict__ out32, double* __restrict__ out64) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * NC) return;
  const int b = i / NC;
  M4 t = from_pose(lidar_sensor + 7 * b);                       // lidar -> ego(t_lidar)
  if (lidar_ego) {
    t = mul(from_pose(lidar_ego + 7 * b), t);                   // -> global
    t = mul(rigid_inverse(from_pose(cam_ego + 7 * i)), t);      // -> ego(t_cam)
  }
  t = mul(rigid_inverse(from_pose(cam_sensor + 7 * i)), t);     // -> camera
  store16(t, out32 ? out32 + 16 * (int64_t)i : nullptr, out64 ? out64 + 16 * (int64_t)i : nullptr);
}

__global__ void compose_sweep2ref_kernel(const double* __restrict__ sweep_sensor, const double* __restrict__ sweep_ego,
                                         const double* __restrict__ ref_sensor, const double* __restrict__ ref_ego, int S,
                                         double* __restrict__ out64) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  M4 t = mul(from_pose(sweep_ego + 7 * s), from_pose(sweep_sensor + 7 * s));   // global_from_car . car_from_current
  t = mul(rigid_inverse(from_pose(ref_ego)), t);                               // car_from_global
  t = mul(rigid_inverse(from_pose(ref_sensor)), t);                            // ref_from_car
  store16(t, nullptr, out64 + 16 * (int64_t)s);
}

// One thread per input point slot (s, p): 16-byte load, transform in double, 16-byte store at the packed position.
__global__ void __launch_bounds__(256)
accumulate_sweeps_kernel(const float4* __restrict__ pts, int64_t sweep_stride_pts, const int32_t* __restrict__ counts,
                         const double* __restrict__ T, int S, int P, float4* __restrict__ out, int32_t* __restrict__ out_total) {
  __shared__ int s_off[65];
  if (threadIdx.x == 0) {
    int run = 0;
    for (int s = 0; s < S; ++s) { s_off[s] = run; run += min(max(counts[s], 0), P); }
    s_off[S] = run;
    if (blockIdx.x == 0 && blockIdx.y == 0 && out_total) *out_total = run;
  }
  __syncthreads();
  const int s = blockIdx.y;
  const int n = s_off[s + 1] - s_off[s];
  const double* t = T + 16 * (int64_t)s;
  double m[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) m[k] = __ldg(t + k);
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
    const float4 q = ld_stream_f4(pts + s * sweep_stride_pts + p);
    const double x = q.x, y = q.y, z = q.z;
    // numpy: trans_matrix.dot(vstack(xyz, 1)) in float64, stored back into the fp32 cloud
    float4 o;
    o.x = (float)fma(m[2], z, fma(m[1], y, fma(m[0], x, m[3])));
    o.y = (float)fma(m[6], z, fma(m[5], y, fma(m[4], x, m[7])));
    o.z = (float)fma(m[10], z, fma(m[9], y, fma(m[8], x, m[11])));
    o.w = q.w;
    st_stream_f4(out + s_off[s] + p, o);
  }
}

}  // namespace calib
}  // namespace tdssf

extern "C" int tdssf_compose_lidar2cam(const double* lidar_sensor, const double* cam_sensor, const double* lidar_ego,
                                       const double* cam_ego, int B, int NC, float* Rt_f32, double* Rt_f64, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(B >= 0 && NC >= 1 && NC <= TDSSF_MAX_CAMS, "tdssf_compose_lidar2cam: B=%d NC=%d out of range", B, NC);
  TDSSF_REQUIRE(lidar_sensor && cam_sensor && (Rt_f32 || Rt_f64), "tdssf_compose_lidar2cam: NULL pointer");
  TDSSF_REQUIRE((lidar_ego == nullptr) == (cam_ego == nullptr), "tdssf_compose_lidar2cam: lidar_ego and cam_ego go together");
  if (B == 0) return TDSSF_OK;
  const int n = B * NC;
  calib::compose_lidar2cam_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(lidar_sensor, cam_sensor, lidar_ego, cam_ego,
                                                                                     B, NC, Rt_f32, Rt_f64);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

extern "C" int tdssf_compose_sweep2ref(const double* sweep_sensor, const double* sweep_ego, const double* ref_sensor,
                                       const double* ref_ego, int S, double* T_f64, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(S >= 0 && S <= 64, "tdssf_compose_sweep2ref: S=%d out of range (0..64)", S);
  TDSSF_REQUIRE(sweep_sensor && sweep_ego && ref_sensor && ref_ego && T_f64, "tdssf_compose_sweep2ref: NULL pointer");
  if (S == 0) return TDSSF_OK;
  calib::compose_sweep2ref_kernel<<<1, 64, 0, (cudaStream_t)stream>>>(sweep_sensor, sweep_ego, ref_sensor, ref_ego, S, T_f64);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

extern "C" int tdssf_accumulate_sweeps(const float* pts, int64_t sweep_stride_floats, const int32_t* counts, const double* T_f64,
                                       int S, int P, float* out, int32_t* out_total, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(S >= 0 && S <= 64 && P >= 0, "tdssf_accumulate_sweeps: S=%d P=%d out of range", S, P);
  TDSSF_REQUIRE((int64_t)S * P < ((int64_t)1 << 31), "tdssf_accumulate_sweeps: S*P too large");
  if (S == 0 || P == 0) {
    if (out_total) TDSSF_CUDA(cudaMemsetAsync(out_total, 0, sizeof(int32_t), (cudaStream_t)stream));
    return TDSSF_OK;
  }
  TDSSF_REQUIRE(pts && counts && T_f64 && out, "tdssf_accumulate_sweeps: NULL pointer");
  TDSSF_REQUIRE(((uintptr_t)pts & 15) == 0 && ((uintptr_t)out & 15) == 0 && sweep_stride_floats % 4 == 0 &&
                    sweep_stride_floats >= 4 * (int64_t)P,
                "tdssf_accumulate_sweeps: points must be (x,y,z,intensity) fp32 rows, 16-byte aligned");
  const unsigned gx = (unsigned)std::min<int64_t>(ceil_div(P, 256), 4 * (int64_t)num_sms());
  calib::accumulate_sweeps_kernel<<<dim3(gx, (unsigned)S), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(pts), sweep_stride_floats / 4, counts, T_f64, S, P, reinterpret_cast<float4*>(out), out_total);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}
