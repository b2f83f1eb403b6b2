// Mean all-reduce of the fusion-MLP gradients over NVLink peer memory: ONE kernel per training step instead of a library
// collective.  (The reference is single-process and has no counterpart; SURVEY §8e defines the semantics: the flat fp32
// gradient of mlp.{0,2,4} - 132 432 values, 0.5 MB - summed over ranks and divided by the world size.)
//
// Every rank owns an exchange buffer (cudaMalloc, exported with cudaIpcGetMemHandle, opened by its peers):
//     flags  [2 parities][kMaxRanks][kMaxChunks] uint32     then     data [2 parities][capacity] float
// One launch, CTA j of rank r:
//   1. copies chunk j of its gradient into its own data[parity], fences, and writes `epoch` into flags[parity][r][j] of
//      EVERY rank's buffer (remote stores over NVLink / NVSwitch);
//   2. waits until flags[parity][p][j] of its OWN buffer (local memory, written remotely) have reached `epoch` for all p;
//   3. reads chunk j of every rank's data[parity] (remote loads), adds them in rank order - every rank computes the same
//      sum, bit for bit - scales by 1 / world and writes the result over its gradient.
// Parities alternate by epoch, so no closing barrier is needed: nobody can write its data for epoch e + 2 before everyone
// has passed step 2 of epoch e + 1, which they only do after finishing the reads of epoch e.  A CTA waits only for data
// that peer CTAs publish before they wait themselves: no co-residency requirement, no deadlock by construction.
#include <algorithm>
#include "common.cuh"

namespace tdssf {
namespace p2p {

constexpr int kMaxRanks = 16, kMaxChunks = 256, kThreads = 256;
constexpr size_t kFlagBytes = (size_t)2 * kMaxRanks * kMaxChunks * sizeof(uint32_t);   // 32 KB

struct Peers { void* buf[kMaxRanks]; };

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {   // coherent at system scope: never from a stale L1 line
  float4 r;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}

__global__ void __launch_bounds__(kThreads)
allreduce_mean_kernel(const Peers peers, int world, int rank, float* __restrict__ grad, int64_t n4, int64_t capacity, uint32_t epoch,
                      unsigned int* debug) {
  const int parity = (int)(epoch & 1u), j = blockIdx.x, chunks = gridDim.x;
  const int64_t per = (n4 + chunks - 1) / chunks, lo = (int64_t)j * per, hi = min(n4, lo + per);
  char* const own = reinterpret_cast<char*>(peers.buf[rank]);
  float4* const own_data = reinterpret_cast<float4*>(own + kFlagBytes) + (int64_t)parity * (capacity / 4);
  float4* const g4 = reinterpret_cast<float4*>(grad);
  // 1. publish this rank's chunk
  for (int64_t i = lo + threadIdx.x; i < hi; i += kThreads) own_data[i] = g4[i];
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < world) {
    uint32_t* f = reinterpret_cast<uint32_t*>(peers.buf[threadIdx.x]) + ((size_t)parity * kMaxRanks + rank) * kMaxChunks + j;
    st_release_sys(f, epoch);
  }
  // 2. wait for every rank's chunk j of this epoch
  if ((int)threadIdx.x < world) {
    const uint32_t* f = reinterpret_cast<const uint32_t*>(own) + ((size_t)parity * kMaxRanks + threadIdx.x) * kMaxChunks + j;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      if (clock64() - t0 > 240000000000LL) {   // ~2 min: a rank never launched its side (ranks may be seconds apart: data loading)
        if (debug) atomicExch(debug, 31u | 0x80000000u);
        __trap();
      }
    }
  }
  __syncthreads();
  // 3. the same sum on every rank: rank order
  const float scale = 1.f / (float)world;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kThreads) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p = 0; p < world; ++p) {
      const float4* d = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(peers.buf[p]) + kFlagBytes) + (int64_t)parity * (capacity / 4);
      const float4 v = ld_peer_f4(d + i);
      acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
    }
    g4[i] = make_float4(acc.x * scale, acc.y * scale, acc.z * scale, acc.w * scale);
  }
}

}  // namespace p2p
}  // namespace tdssf

extern "C" size_t tdssf_p2p_buffer_bytes(int64_t max_floats) {
  const size_t cap = ((size_t)std::max<int64_t>(max_floats, 4) + 3) & ~(size_t)3;
  return tdssf::p2p::kFlagBytes + 2 * cap * sizeof(float);
}

// Allocates and zeroes this rank's exchange buffer and exports it: `handle_out` receives 64 bytes (cudaIpcMemHandle_t) for
// the peers.  One-time set-up (this call allocates and synchronises; the per-step call below does neither).
extern "C" int tdssf_p2p_create(int64_t max_floats, void** buf_out, void* handle_out_host) {
  using namespace tdssf;
  TDSSF_REQUIRE(buf_out && handle_out_host && max_floats > 0, "tdssf_p2p_create: bad arguments");
  const size_t bytes = tdssf_p2p_buffer_bytes(max_floats);
  void* buf = nullptr;
  TDSSF_CUDA(cudaMalloc(&buf, bytes));
  TDSSF_CUDA(cudaMemset(buf, 0, bytes));
  TDSSF_CUDA(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  const cudaError_t e = cudaIpcGetMemHandle(&h, buf);
  if (e != cudaSuccess) {
    cudaFree(buf);
    set_error("tdssf_p2p_create: cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    cudaGetLastError();
    return TDSSF_ENOSUP;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  memcpy(handle_out_host, &h, 64);
  *buf_out = buf;
  return TDSSF_OK;
}

// Maps a peer's exchange buffer (the 64 bytes its tdssf_p2p_create exported) into this process.
extern "C" int tdssf_p2p_open(const void* handle_host, void** peer_buf_out) {
  using namespace tdssf;
  TDSSF_REQUIRE(handle_host && peer_buf_out, "tdssf_p2p_open: NULL pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle_host, 64);
  void* p = nullptr;
  const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    set_error("tdssf_p2p_open: cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e));
    cudaGetLastError();
    return TDSSF_ENOSUP;
  }
  *peer_buf_out = p;
  return TDSSF_OK;
}

extern "C" int tdssf_p2p_close(void* peer_buf) {
  if (peer_buf) cudaIpcCloseMemHandle(peer_buf);
  return TDSSF_OK;
}
extern "C" int tdssf_p2p_destroy(void* own_buf) {
  if (own_buf) cudaFree(own_buf);
  return TDSSF_OK;
}

// grad[0..n_floats) := mean over ranks, in place, one launch on `stream`.  bufs_host: world pointers (this process's
// mappings of every rank's exchange buffer, own buffer at index rank); `epoch` counts this group's calls from 1 and must
// be the same on every rank; max_floats is what the buffers were created with.
extern "C" int tdssf_p2p_allreduce_mean(void* const* bufs_host, int world, int rank, float* grad, int64_t n_floats, int64_t max_floats,
                                        uint32_t epoch, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(bufs_host && grad, "tdssf_p2p_allreduce_mean: NULL pointer");
  TDSSF_REQUIRE(world >= 1 && world <= p2p::kMaxRanks && rank >= 0 && rank < world, "tdssf_p2p_allreduce_mean: bad rank %d / world %d (max %d)", rank, world,
                p2p::kMaxRanks);
  TDSSF_REQUIRE(n_floats >= 0 && n_floats <= max_floats && (n_floats & 3) == 0 && ((uintptr_t)grad & 15) == 0,
                "tdssf_p2p_allreduce_mean: n_floats must be a multiple of 4 within the buffers' capacity, grad 16-byte aligned");
  TDSSF_REQUIRE(epoch != 0, "tdssf_p2p_allreduce_mean: epochs count from 1");
  if (n_floats == 0) return TDSSF_OK;
  p2p::Peers peers = {};
  for (int p = 0; p < world; ++p) {
    TDSSF_REQUIRE(bufs_host[p] != nullptr, "tdssf_p2p_allreduce_mean: NULL buffer of rank %d", p);
    peers.buf[p] = bufs_host[p];
  }
  const int64_t n4 = n_floats / 4;
  const int64_t cap = (std::max<int64_t>(max_floats, 4) + 3) & ~(int64_t)3;
  const int chunks = (int)std::min<int64_t>(p2p::kMaxChunks, std::max<int64_t>(1, std::min<int64_t>(num_sms(), ceil_div(n4, p2p::kThreads))));
  p2p::allreduce_mean_kernel<<<chunks, p2p::kThreads, 0, (cudaStream_t)stream>>>(peers, world, rank, grad, n4, cap, epoch, nullptr);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}
