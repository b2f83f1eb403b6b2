// a7: devoxelize (utils/lidar.py:120-172): nearest voxel centre per raw point, gather of the voxel
// scores, and the backward scatter-add.
//
// Nearest-voxel search, exact brute force (the parity mode): every CTA owns kPts points (one per
// thread, 2 per thread via ILP) and streams the sweep's voxel centres through shared memory in
// tiles; distance = dx*dx + dy*dy + dz*dz in fp32 (direct differences — more accurate than the
// reference's cdist matmul expansion), strict '<' so the first index wins ties.
// O(P*V) fp32 CUDA-core work: compute-bound, reported separately from the HBM-bound kernels.
//
// Grid-accelerated search with the SAME result, bit for bit (tdssf_devox_nn with a workspace): the sweep's voxel
// centres are linked into a hash table of cubic cells of edge `cell`; a point evaluates the voxels of the 27 cells
// around it with the brute-force distance arithmetic and tie rule (lowest index among equal distances).  Every voxel
// closer than `cell` lies in those cells, so a best distance below (0.99 cell)^2 is the brute-force answer; the other
// points (sparse regions, non-finite coordinates) go on a list and a warp per point scans all voxels.
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

// ---- grid-accelerated exact search ------------------------------------------------------------------------------
__device__ __forceinline__ float dist2_rn(float px, float py, float pz, float cx, float cy, float cz) {
  const float dx = px - cx, dy = py - cy, dz = pz - cz;
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
__device__ __forceinline__ uint32_t cell_hash(int ix, int iy, int iz, uint32_t mask) {
  return (((uint32_t)ix * 73856093u) ^ ((uint32_t)iy * 19349663u) ^ ((uint32_t)iz * 83492791u)) & mask;
}
__device__ __forceinline__ int cell_of(float x, float inv_cell) { return __float2int_rd(__fmul_rn(x, inv_cell)); }

// one thread per voxel: node = (x, y, z, next voxel of the same bucket), pushed at the head of its bucket's list.
// Rows at exactly (0,0,0) - the zero padding of voxel_raw up to V_max, hundreds to thousands of identical
// rows - are not linked (they would form one long serial bucket walk for every point near the origin): only the lowest
// such index can ever win (equal distance, first index wins), so it is kept in first_zero[b] and offered to every
// query as one extra candidate.
__global__ void devox_grid_build_kernel(const float* __restrict__ vox, int64_t vox_stride, int V, float inv_cell,
                                        int* __restrict__ head, uint32_t mask, float4* __restrict__ node,
                                        int* __restrict__ first_zero) {
  const int b = blockIdx.y;
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= V) return;
  const float* q = vox + ((int64_t)b * V + v) * vox_stride;
  const float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
  if (x == 0.f && y == 0.f && z == 0.f) {
    atomicMin(first_zero + b, v);
    return;
  }
  const uint32_t h = cell_hash(cell_of(x, inv_cell), cell_of(y, inv_cell), cell_of(z, inv_cell), mask);
  const int old = atomicExch(head + (int64_t)b * (mask + 1) + h, v);
  node[(int64_t)b * V + v] = make_float4(x, y, z, __int_as_float(old));
}

// a warp per point: lane l < 27 walks the bucket of cell l of the 3x3x3 neighbourhood, then the warp reduces to the
// smallest distance (lowest index among equals); unresolved points are appended to `todo`
constexpr int kQueryThreads = 256;
__global__ void __launch_bounds__(kQueryThreads)
devox_grid_query_kernel(const float* __restrict__ pts, int64_t pts_stride, int P, int V, float inv_cell, float accept_d2,
                        const int* __restrict__ head, uint32_t mask, const float4* __restrict__ node,
                        const int* __restrict__ first_zero, int64_t* __restrict__ idx, int* __restrict__ todo,
                        int* __restrict__ n_todo) {
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  const int vz = __ldg(first_zero + b);  // lowest row at the origin, or >= V when there is none
  const int* hb = head + (int64_t)b * (mask + 1);
  const float4* nb = node + (int64_t)b * V;
  const int dx = lane % 3 - 1, dy = (lane / 3) % 3 - 1, dz = lane / 9 - 1;  // lanes 27..31: no cell
  for (int p = (blockIdx.x * kQueryThreads + threadIdx.x) >> 5; p < P; p += (gridDim.x * kQueryThreads) >> 5) {
    const float* q = pts + ((int64_t)b * P + p) * pts_stride;
    const float px = __ldg(q), py = __ldg(q + 1), pz = __ldg(q + 2);
    float best = INFINITY;
    int besti = 0x7fffffff;
    if (lane < 27) {
      int v = __ldg(hb + cell_hash(cell_of(px, inv_cell) + dx, cell_of(py, inv_cell) + dy, cell_of(pz, inv_cell) + dz, mask));
      while (v >= 0) {
        const float4 c = __ldg(nb + v);
        const float d = dist2_rn(px, py, pz, c.x, c.y, c.z);
        if (d < best || (d == best && v < besti)) { best = d; besti = v; }
        v = __float_as_int(c.w);
      }
    } else if (lane == 27 && vz < V) {  // the origin rows' representative
      best = dist2_rn(px, py, pz, 0.f, 0.f, 0.f);
      besti = vz;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
      if (ob < best || (ob == best && oi < besti)) { best = ob; besti = oi; }
    }
    if (lane == 0) {
      if (best <= accept_d2)  // false for NaN / infinite distances too
        idx[(int64_t)b * P + p] = besti;
      else
        todo[atomicAdd(n_todo, 1)] = b * P + p;
    }
  }
}

// a block per unresolved point: all V voxels, four loads in flight per thread; lowest index among equal distances
// (== the sequential strict '<' scan of the brute-force kernel)
constexpr int kScanThreads = 256;
__global__ void __launch_bounds__(kScanThreads)
devox_scan_todo_kernel(const float* __restrict__ pts, int64_t pts_stride, const float* __restrict__ vox, int64_t vox_stride,
                       int P, int V, const int* __restrict__ todo, const int* __restrict__ n_todo, int64_t* __restrict__ idx) {
  __shared__ float s_best[kScanThreads / 32];
  __shared__ int s_idx[kScanThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = *n_todo;
  for (int w = blockIdx.x; w < n; w += gridDim.x) {
    const int gp = todo[w], b = gp / P;
    const float* q = pts + (int64_t)gp * pts_stride;
    const float px = __ldg(q), py = __ldg(q + 1), pz = __ldg(q + 2);
    const float* vb = vox + (int64_t)b * V * vox_stride;
    float best = INFINITY;
    int besti = 0x7fffffff;
    for (int v0 = threadIdx.x; v0 < V; v0 += 4 * kScanThreads) {
      float d[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int v = min(v0 + k * kScanThreads, V - 1);
        const float* c = vb + (int64_t)v * vox_stride;
        d[k] = dist2_rn(px, py, pz, __ldg(c), __ldg(c + 1), __ldg(c + 2));
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int v = v0 + k * kScanThreads;   // ascending per thread: strict '<' keeps the lowest index
        if (v < V && d[k] < best) { best = d[k]; besti = v; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
      if (ob < best || (ob == best && oi < besti)) { best = ob; besti = oi; }
    }
    if (lane == 0) { s_best[warp] = best; s_idx[warp] = besti; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int k = 1; k < kScanThreads / 32; ++k)
        if (s_best[k] < best || (s_best[k] == best && s_idx[k] < besti)) { best = s_best[k]; besti = s_idx[k]; }
      // no finite distance at all (NaN point, NaN voxels): the sequential scan never updates its initial index 0
      idx[gp] = besti == 0x7fffffff ? 0 : besti;
    }
    __syncthreads();
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

static uint32_t devox_table_size(int V) {
  uint32_t t = 1024;
  while (t < 2u * (uint32_t)V) t <<= 1;
  return t;
}
// workspace: [B][T] bucket heads, [B][V] nodes (16 B), [B*P] unresolved points, counter, [B] first origin row
static size_t devox_ws_bytes(int B, int P, int V) {
  const size_t T = devox_table_size(V);
  return (((size_t)B * T * 4 + 255) & ~(size_t)255) + (((size_t)B * V * 16 + 255) & ~(size_t)255) +
         (((size_t)B * P * 4 + 255) & ~(size_t)255) + 256 + (((size_t)B * 4 + 255) & ~(size_t)255);
}

extern "C" size_t tdssf_devox_nn_workspace_bytes(int B, int P, int V) {
  if (B <= 0 || P <= 0 || V <= 0) return 256;
  return devox_ws_bytes(B, P, V);
}

extern "C" int tdssf_devox_nn(const float* pts, int64_t pts_stride, const float* vox, int64_t vox_stride,
                              int B, int P, int V, int64_t* idx, float cell, void* workspace, size_t workspace_bytes,
                              void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(pts && vox && idx, "tdssf_devox_nn: NULL pointer");
  TDSSF_REQUIRE(pts_stride >= 3 && vox_stride >= 3, "tdssf_devox_nn: strides must be >= 3 floats");
  TDSSF_REQUIRE(B >= 0 && P >= 0 && V >= 1, "tdssf_devox_nn: need V >= 1 (argmin over an empty set)");
  if (B == 0 || P == 0) return TDSSF_OK;
  cudaStream_t s = (cudaStream_t)stream;
  if (!workspace) {  // exact brute force
    dim3 grid((unsigned)ceil_div(P, kNNThreads * kNNPtsPerThread), (unsigned)B);
    devox_nn_kernel<<<grid, kNNThreads, 0, s>>>(pts, pts_stride, vox, vox_stride, P, V, idx);
    TDSSF_LAUNCH_CHECK();
    return TDSSF_OK;
  }
  TDSSF_REQUIRE(cell > 0.f && cell < 1e30f, "tdssf_devox_nn: cell edge must be a positive finite length, got %g", (double)cell);
  TDSSF_REQUIRE((int64_t)B * P < (1ll << 31), "tdssf_devox_nn: B*P exceeds the 32-bit index range");
  TDSSF_REQUIRE(((uintptr_t)workspace & 255) == 0 && workspace_bytes >= devox_ws_bytes(B, P, V),
                "tdssf_devox_nn: 256-byte aligned workspace of %zu bytes needed (tdssf_devox_nn_workspace_bytes), got %zu",
                devox_ws_bytes(B, P, V), workspace_bytes);
  const uint32_t T = devox_table_size(V);
  char* w = reinterpret_cast<char*>(workspace);
  int* head = reinterpret_cast<int*>(w);
  w += ((size_t)B * T * 4 + 255) & ~(size_t)255;
  float4* node = reinterpret_cast<float4*>(w);
  w += ((size_t)B * V * 16 + 255) & ~(size_t)255;
  int* todo = reinterpret_cast<int*>(w);
  w += ((size_t)B * P * 4 + 255) & ~(size_t)255;
  int* n_todo = reinterpret_cast<int*>(w);
  w += 256;
  int* first_zero = reinterpret_cast<int*>(w);
  TDSSF_CUDA(cudaMemsetAsync(head, 0xff, (size_t)B * T * 4, s));  // -1: empty bucket
  TDSSF_CUDA(cudaMemsetAsync(n_todo, 0, 4, s));
  TDSSF_CUDA(cudaMemsetAsync(first_zero, 0x7f, (size_t)B * 4, s));  // 0x7f7f7f7f: no origin row
  const float inv_cell = 1.f / cell;
  const float accept = 0.99f * cell;
  devox_grid_build_kernel<<<dim3((unsigned)ceil_div(V, 256), (unsigned)B), 256, 0, s>>>(vox, vox_stride, V, inv_cell, head,
                                                                                   T - 1, node, first_zero);
  TDSSF_LAUNCH_CHECK();
  const unsigned qblocks = (unsigned)std::min<int64_t>(ceil_div((int64_t)P * 32, kQueryThreads), (int64_t)num_sms() * 16);
  devox_grid_query_kernel<<<dim3(qblocks, (unsigned)B), kQueryThreads, 0, s>>>(pts, pts_stride, P, V, inv_cell, accept * accept,
                                                                              head, T - 1, node, first_zero, idx, todo, n_todo);
  TDSSF_LAUNCH_CHECK();
  devox_scan_todo_kernel<<<(unsigned)num_sms() * 8, kScanThreads, 0, s>>>(pts, pts_stride, vox, vox_stride, P, V, todo, n_todo,
                                                                         idx);
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

// ---- tdssf_devox_nn_static: the grid-accelerated search for callers that cannot allocate device memory ------------
// No counterpart in the reference.  The hash-grid workspace lives in a
// file-scope device array (one instance per GPU that loads the library, held until exit) instead of a caller-owned
// buffer.  In host code `g_devox_static_ws` is only a handle: its device address is fetched once per device with
// cudaGetSymbolAddress while that device is current, and that address is what the kernels receive.  Every call shares
// the array (the bucket heads, the unresolved-point counter and the origin rows are re-initialised by every call, so
// nothing of the previous call is read); calls that overlap on the GPU race on it.
constexpr size_t kDevoxStaticBytes = (size_t)TDSSF_DEVOX_NN_STATIC_BYTES;
static __device__ __align__(256) unsigned char g_devox_static_ws[kDevoxStaticBytes];

static int devox_static_ws(void** out) {
  static void* base[tdssf::kMaxDevices] = {};
  const int dev = tdssf::current_device();
  if (!base[dev]) {
    void* p = nullptr;
    TDSSF_CUDA(cudaGetSymbolAddress(&p, g_devox_static_ws));
    TDSSF_REQUIRE(p && ((uintptr_t)p & 255) == 0, "tdssf_devox_nn_static: bad address %p for the static workspace", p);
    base[dev] = p;
  }
  *out = base[dev];
  return TDSSF_OK;
}

extern "C" int tdssf_devox_nn_static(const float* pts, int64_t pts_stride, const float* vox, int64_t vox_stride,
                                     int B, int P, int V, int64_t* idx, float cell, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(B >= 0 && P >= 0 && V >= 1, "tdssf_devox_nn_static: need B >= 0, P >= 0, V >= 1");
  if (B == 0 || P == 0) return TDSSF_OK;
  const size_t need = devox_ws_bytes(B, P, V);
  TDSSF_REQUIRE(need <= kDevoxStaticBytes,
                "tdssf_devox_nn_static: B=%d P=%d V=%d needs %zu workspace bytes, the static workspace holds %zu "
                "(TDSSF_DEVOX_NN_STATIC_BYTES); use tdssf_devox_nn with a caller-owned workspace",
                B, P, V, need, kDevoxStaticBytes);
  void* ws = nullptr;
  const int rc = devox_static_ws(&ws);
  if (rc != TDSSF_OK) return rc;
  return tdssf_devox_nn(pts, pts_stride, vox, vox_stride, B, P, V, idx, cell, ws, kDevoxStaticBytes, stream);
}
