// Micro-test: does cp.async.bulk.tensor.2d ... tile::gather4 deliver four arbitrary rows of a bf16 [rows, 384] matrix
// into a 128B-swizzled K-major slab the way the fused kernel's UMMA descriptors expect (row r at r*128, 16-byte chunk c
// at ((c ^ (r & 7)) << 4))?   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/micro/gather4_probe.bin tools/micro/gather4_probe.cu
// Prints one JSON line per tensor-map variant (box rows 1 or 4): mismatches against the expected slab.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1)
gather_kernel(const __grid_constant__ CUtensorMap tm, const int* __restrict__ rows, int col0, uint16_t* __restrict__ out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(16384) : "memory");
    for (int g = 0; g < 32; ++g) {
      const int r0 = rows[4 * g], r1 = rows[4 * g + 1], r2 = rows[4 * g + 2], r3 = rows[4 * g + 3];
      asm volatile(
          "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(
              smem_u32(smem + g * 512)),
          "l"(&tm), "r"(col0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_u32(bar))
          : "memory");
    }
  }
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(0) : "memory");
  // de-swizzle: row r = thread, 8 chunks of 16 bytes
  const int r = threadIdx.x;
  for (int c = 0; c < 8; ++c) {
    const uint4 v = *reinterpret_cast<const uint4*>(smem + r * 128 + ((c ^ (r & 7)) << 4));
    *reinterpret_cast<uint4*>(out + r * 64 + c * 8) = v;
  }
}


// Throughput: every CTA gathers `n_slabs` slabs of 128 random rows x 64 channels through a ring of kRing 16 KB stages,
// `issuers` threads issuing (each a share of the 32 gather4 of a slab).  Reports bytes landed per second.
template <int kRing>
__global__ void __launch_bounds__(128, 1)
gather_rate_kernel(const __grid_constant__ CUtensorMap tm, const int* __restrict__ rows, int n_rows_list, int n_slabs, int issuers,
                   unsigned* sink) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kRing * 16384);
  if (threadIdx.x == 0) {
    for (int s = 0; s < kRing; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bars + s)), "r"(issuers));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if ((int)threadIdx.x >= issuers) return;
  const int per = 32 / issuers;
  const int* my = rows + ((size_t)blockIdx.x * 977 % 1000) * 128;
  for (int it = 0; it < n_slabs + kRing; ++it) {
    if (it >= kRing) {   // wait for the slab issued kRing iterations ago (frees its stage)
      const int d = it - kRing, s = d % kRing;
      const uint32_t parity = (d / kRing) & 1;
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bars + s)), "r"(parity) : "memory");
    }
    if (it < n_slabs) {
      const int s = it % kRing;
      const int* rr = my + ((size_t)it * 128) % (size_t)(n_rows_list - 128 - 1000 * 128 > 0 ? n_rows_list - 1000 * 128 - 128 : 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bars + s)), "r"(per * 512) : "memory");
      for (int g = threadIdx.x * per; g < (int)(threadIdx.x + 1) * per; ++g) {
        const int r0 = rr[4 * g], r1 = rr[4 * g + 1], r2 = rr[4 * g + 2], r3 = rr[4 * g + 3];
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(
                smem_u32(smem + s * 16384 + g * 512)),
            "l"(&tm), "r"((it % 6) * 64), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_u32(bars + s))
            : "memory");
      }
    }
  }
  if (smem[0] == 0x5a && smem[1] == 0xa5 && smem[2] == 0x77 && smem[3] == 1) *sink = 1;
}

int main() {
  const int R = 20000, C = 384, col0 = 128;
  std::vector<uint16_t> h((size_t)R * C);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (uint16_t)((i * 2654435761u) >> 13);
  std::vector<int> rows(128);
  for (int i = 0; i < 128; ++i) rows[i] = (int)((i * 7919u + 13u * i * i) % R);
  rows[5] = rows[4];   // a repeated row
  uint16_t *d, *dout;
  int* drows;
  cudaMalloc(&d, h.size() * 2);
  cudaMalloc(&dout, 128 * 64 * 2);
  cudaMalloc(&drows, 128 * 4);
  cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(drows, rows.data(), 128 * 4, cudaMemcpyHostToDevice);
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
    printf("{\"error\": \"no cuTensorMapEncodeTiled\"}\n");
    return 1;
  }
  EncodeTiledFn enc = reinterpret_cast<EncodeTiledFn>(fp);
  cudaFuncSetAttribute(gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 + 1024 + 64);
  for (int box_rows : {1}) {   // (box rows 4: the instruction faults - a gather4 map has one-row boxes)
    CUtensorMap tm;
    const cuuint64_t dims[2] = {(cuuint64_t)C, (cuuint64_t)R};
    const cuuint64_t strides[1] = {(cuuint64_t)C * 2};
    const cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult rc = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) {
      printf("{\"box_rows\": %d, \"encode_rc\": %d}\n", box_rows, (int)rc);
      continue;
    }
    cudaMemset(dout, 0xff, 128 * 64 * 2);
    gather_kernel<<<1, 128, 16384 + 1024 + 64>>>(tm, drows, col0, dout);
    const cudaError_t e = cudaDeviceSynchronize();
    std::vector<uint16_t> o(128 * 64);
    cudaMemcpy(o.data(), dout, o.size() * 2, cudaMemcpyDeviceToHost);
    int bad = 0, first = -1;
    for (int r = 0; r < 128; ++r)
      for (int c = 0; c < 64; ++c)
        if (o[r * 64 + c] != h[(size_t)rows[r] * C + col0 + c]) { if (first < 0) first = r * 64 + c; ++bad; }
    printf("{\"box_rows\": %d, \"cuda\": \"%s\", \"mismatches\": %d, \"first\": %d}\n", box_rows, cudaGetErrorString(e), bad, first);
    if (e != cudaSuccess) break;   // a sticky error: the context is gone
  }

  {  // throughput over a config-2 sized token set: 8 sweeps x 6 cameras x 2840 rows of 384 bf16
    const int RT = 8 * 6 * 2840, LIST = 1 << 20;
    uint16_t* dt; int* dl; unsigned* sink;
    cudaMalloc(&dt, (size_t)RT * C * 2); cudaMemset(dt, 1, (size_t)RT * C * 2);
    std::vector<int> lst(LIST);
    uint32_t x = 12345u;
    for (int i = 0; i < LIST; ++i) { x = x * 1664525u + 1013904223u; lst[i] = (int)((x >> 8) % RT); }
    cudaMalloc(&dl, LIST * 4); cudaMemcpy(dl, lst.data(), LIST * 4, cudaMemcpyHostToDevice);
    cudaMalloc(&sink, 4);
    CUtensorMap tm;
    const cuuint64_t dims[2] = {(cuuint64_t)C, (cuuint64_t)RT};
    const cuuint64_t strides[1] = {(cuuint64_t)C * 2};
    const cuuint32_t box[2] = {64, 1};
    const cuuint32_t estr[2] = {1, 1};
    enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dt, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int n_slabs = 6 * 15;   // the token slabs of 15 tiles
    auto run = [&](auto kern, int ring, int issuers) {
      const int sm = ring * 16384 + 1024 + 128;
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      float best = 1e9f;
      for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        kern<<<sms, 128, sm>>>(tm, dl, LIST, n_slabs, issuers, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
      }
      printf("{\"ring\": %d, \"issuers\": %d, \"slabs_per_cta\": %d, \"us\": %.1f, \"GBps\": %.0f, \"err\": \"%s\"}\n", ring, issuers, n_slabs,
             best * 1e3, (double)sms * n_slabs * 16384 / (best * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
    };
    run(gather_rate_kernel<3>, 3, 1); run(gather_rate_kernel<3>, 3, 4); run(gather_rate_kernel<3>, 3, 32);
    run(gather_rate_kernel<6>, 6, 1); run(gather_rate_kernel<6>, 6, 4); run(gather_rate_kernel<6>, 6, 32);
    run(gather_rate_kernel<10>, 10, 8);
  }
  return 0;
}
