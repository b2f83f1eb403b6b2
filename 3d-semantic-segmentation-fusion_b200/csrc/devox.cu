// a7: devoxelize (utils/lidar.py:120-172): nearest voxel centre per raw point, gather of the voxel
// scores, and the backward scatter-add.
//
// Nearest-voxel search, exact brute force (the parity mode): every CTA owns kPts points (one per
// thread, 2 per thread via ILP) and streams the sweep's voxel centres through shared memory in
// tiles; distance = dx*dx + dy*dy + dz*dz in fp32 (direct differences — more accurate than the
// reference's cdist matmul expansion), strict '<' so the first index wins ties.
// O(P*V) fp32 CUDA-core work: compute-bound, reported separately from the HBM-bound kernels.
#include <algorithm>
#include "common.cuh"

namespace tdssf {

constexpr int kNNThreads = 128;
constexpr int kNNPtsPerThread = 2;
constexpr int kNNTile = 1024;  // voxels staged per iteration (16 KB)

__global__ void __launch_bounds__(kNNThreads)
devox_nn_kernel(const float* __restrict__ pts, int64_t pts_stride, const float* __restrict__ vox,
                int64_t vox_stride, int P, int V, int64_t* __restrict__ idx) {
  __shared__ float4 s_vox[kNNTile];
  const int b = blockIdx.y;
  const float* pb = pts + (int64_t)b * P * pts_stride;
  const float* vb = vox + (int64_t)b * V * vox_stride;
  float px[kNNPtsPerThread], py[kNNPtsPerThread], pz[kNNPtsPerThread], best[kNNPtsPerThread];
  int besti[kNNPtsPerThread];
  const int p0 = blockIdx.x * (kNNThreads * kNNPtsPerThread) + threadIdx.x;
#pragma unroll
  for (int i = 0; i < kNNPtsPerThread; ++i) {
    const int p = min(p0 + i * kNNThreads, P - 1);
    px[i] = __ldg(pb + p * pts_stride);
    py[i] = __ldg(pb + p * pts_stride + 1);
    pz[i] = __ldg(pb + p * pts_stride + 2);
    best[i] = INFINITY;
    besti[i] = 0;
  }
  for (int v0 = 0; v0 < V; v0 += kNNTile) {
    const int n = min(kNNTile, V - v0);
    __syncthreads();
    for (int j = threadIdx.x; j < n; j += kNNThreads) {
      const float* q = vb + (int64_t)(v0 + j) * vox_stride;
      s_vox[j] = make_float4(__ldg(q), __ldg(q + 1), __ldg(q + 2), 0.f);
    }
    __syncthreads();
#pragma unroll 4
    for (int j = 0; j < n; ++j) {
      const float4 c = s_vox[j];
#pragma unroll
      for (int i = 0; i < kNNPtsPerThread; ++i) {
        const float dx = px[i] - c.x, dy = py[i] - c.y, dz = pz[i] - c.z;
        const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        if (d < best[i]) { best[i] = d; besti[i] = v0 + j; }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kNNPtsPerThread; ++i) {
    const int p = p0 + i * kNNThreads;
    if (p < P) idx[(int64_t)b * P + p] = besti[i];
  }
}

// out[b,p,:] = scores[b, idx[b,p], :]; one thread per 16-byte vector (C % 4 == 0) or scalar.
__global__ void devox_gather_kernel(const float* __restrict__ scores, const int64_t* __restrict__ idx,
                                    int64_t n_rows, int P, int V, int C, float* __restrict__ out) {
  const int vec = (C % 4 == 0) ? 4 : 1;
  const int per_row = C / vec;
  const int64_t total = n_rows * per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / per_row;
    const int c = (int)(i % per_row) * vec;
    const int64_t b = row / P;
    const int64_t src = (b * V + idx[row]) * C + c;
    if (vec == 4)
      *reinterpret_cast<float4*>(out + row * C + c) = __ldg(reinterpret_cast<const float4*>(scores + src));
    else
      out[row * C + c] = __ldg(scores + src);
  }
}

__global__ void devox_scatter_kernel(const float* __restrict__ dout, const int64_t* __restrict__ idx,
                                     int64_t n_rows, int P, int V, int C, float* __restrict__ dscores) {
  const int64_t total = n_rows * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / C;
    const int c = (int)(i % C);
    const int64_t b = row / P;
    atomicAdd(dscores + (b * V + idx[row]) * C + c, __ldg(dout + i));
  }
}

}  // namespace tdssf

extern "C" int tdssf_devox_nn(const float* pts, int64_t pts_stride, const float* vox, int64_t vox_stride,
                              int B, int P, int V, int64_t* idx, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(pts && vox && idx, "tdssf_devox_nn: NULL pointer");
  TDSSF_REQUIRE(pts_stride >= 3 && vox_stride >= 3, "tdssf_devox_nn: strides must be >= 3 floats");
  TDSSF_REQUIRE(B >= 0 && P >= 0 && V >= 1, "tdssf_devox_nn: need V >= 1 (argmin over an empty set)");
  if (B == 0 || P == 0) return TDSSF_OK;
  dim3 grid((unsigned)ceil_div(P, kNNThreads * kNNPtsPerThread), (unsigned)B);
  devox_nn_kernel<<<grid, kNNThreads, 0, (cudaStream_t)stream>>>(pts, pts_stride, vox, vox_stride, P, V, idx);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

extern "C" int tdssf_devox_gather(const float* scores, const int64_t* idx, int B, int P, int V, int C,
                                  float* out, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(scores && idx && out, "tdssf_devox_gather: NULL pointer");
  TDSSF_REQUIRE(C >= 1 && V >= 1, "tdssf_devox_gather: bad C/V");
  const int64_t n_rows = (int64_t)B * P;
  if (n_rows == 0) return TDSSF_OK;
  if (C % 4 == 0)
    TDSSF_REQUIRE((((uintptr_t)scores | (uintptr_t)out) & 15) == 0, "tdssf_devox_gather: 16-byte alignment required");
  const int64_t work = n_rows * (C % 4 == 0 ? C / 4 : C);
  const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div(work, 256), (int64_t)num_sms() * 32);
  devox_gather_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(scores, idx, n_rows, P, V, C, out);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

extern "C" int tdssf_devox_scatter_add(const float* dout, const int64_t* idx, int B, int P, int V, int C,
                                       float* dscores, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(dout && idx && dscores, "tdssf_devox_scatter_add: NULL pointer");
  TDSSF_REQUIRE(C >= 1 && V >= 1, "tdssf_devox_scatter_add: bad C/V");
  const int64_t n_rows = (int64_t)B * P;
  if (n_rows == 0) return TDSSF_OK;
  const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div(n_rows * C, 256), (int64_t)num_sms() * 32);
  devox_scatter_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dout, idx, n_rows, P, V, C, dscores);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}
