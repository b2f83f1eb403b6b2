// Micro-benchmark: how many bytes per second can the 148 SMs of a B200 pull out of L2 / HBM, by path?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/l2_ingest.bin tools/micro/l2_ingest.cu && tools/micro/l2_ingest.bin
// Reads a buffer of `MB` megabytes `reps` times with every SM (one 1024-thread CTA per SM):
//   path ldg   : ld.global.v4 (L1-allocating), each warp streaming 512-byte pieces, 8 loads in flight per thread
//   path bulk  : cp.async.bulk global -> shared (TMA engine, no registers / LSU), 16 KB copies, 8 in flight per CTA
// With a 48 MB buffer the second and later passes hit in the 126 MB L2 (fabric-bound: L2 -> SM); with 2 GB every pass
// comes from HBM.  The fused fusion kernel moves about 463 MB across that fabric per launch for 303 MB of algorithmic
// (DRAM) bytes, so this number - not the HBM peak - is what bounds it.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(1024, 1) read_ldg(const uint4* __restrict__ buf, size_t n_vec, int reps, unsigned* sink) {
  unsigned acc = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 7 * stride < n_vec; i += 8 * stride) {
      uint4 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[k].x), "=r"(v[k].y), "=r"(v[k].z), "=r"(v[k].w) : "l"(buf + i + k * stride));
#pragma unroll
      for (int k = 0; k < 8; ++k) acc ^= v[k].x ^ v[k].y ^ v[k].z ^ v[k].w;
    }
  }
  if (acc == 0x12345678u) *sink = acc;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1) read_bulk(const uint8_t* __restrict__ buf, size_t bytes, int reps, unsigned* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];   // 8 stages x 16 KB + barriers
  constexpr int kStages = 8, kChunk = 16384;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kChunk);
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + s)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const size_t n_chunks = bytes / kChunk;
  size_t issued = 0, done = 0;
  const size_t total = ((n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x) * (size_t)reps;
  auto src = [&](size_t k) { const size_t per = (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x; return buf + (blockIdx.x + (k % per) * gridDim.x) * (size_t)kChunk; };
  while (done < total) {
    while (issued < total && issued - done < kStages) {
      const int s = issued % kStages;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bars + s)), "r"(kChunk) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem + s * kChunk)),
                   "l"(src(issued)), "r"(kChunk), "r"(smem_u32(bars + s))
                   : "memory");
      ++issued;
    }
    const int s = done % kStages;
    const uint32_t parity = (done / kStages) & 1;
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bars + s)), "r"(parity) : "memory");
    ++done;
  }
  if (smem[0] == 0x5a && smem[1] == 0xa5 && smem[2] == 0x77) *sink = 1;
}

int main(int argc, char** argv) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  unsigned* sink;
  cudaMalloc(&sink, 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int smem_bulk = 8 * 16384 + 64;
  cudaFuncSetAttribute(read_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bulk);
  const size_t sizes_mb[3] = {48, 96, 2048};
  for (size_t mb : sizes_mb) {
    const size_t bytes = mb << 20;
    uint8_t* buf;
    if (cudaMalloc(&buf, bytes) != cudaSuccess) { printf("alloc %zu MB failed\n", mb); continue; }
    cudaMemset(buf, 1, bytes);
    const int reps = mb <= 96 ? 20 : 2;
    for (int path = 0; path < 2; ++path) {
      for (int warm = 0; warm < 2; ++warm) {
        cudaEventRecord(e0);
        if (path == 0) read_ldg<<<sms, 1024>>>(reinterpret_cast<const uint4*>(buf), bytes / 16, reps, sink);
        else read_bulk<<<sms, 128, smem_bulk>>>(buf, bytes, reps, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
      }
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const cudaError_t err = cudaGetLastError();
      printf("{\"path\": \"%s\", \"buffer_MB\": %zu, \"passes\": %d, \"ms\": %.3f, \"GBps\": %.1f, \"sms\": %d, \"err\": \"%s\"}\n", path ? "cp.async.bulk" : "ld.global.v4",
             mb, reps, ms, (double)bytes * reps / (ms * 1e-3) / 1e9, sms, cudaGetErrorString(err));
    }
    cudaFree(buf);
  }
  return 0;
}
