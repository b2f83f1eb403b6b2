// Micro-benchmark: cycles per st.shared instruction for the producers' swizzled-slab store patterns (one warp, then 8 warps).
// nvcc -gencode arch=compute_100a,code=sm_100a -o sts_pattern sts_pattern.cu && ./sts_pattern
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ void sts64(uint32_t a, uint32_t x, uint32_t y) { asm volatile("st.shared.v2.b32 [%0], {%1,%2};" ::"r"(a), "r"(x), "r"(y) : "memory"); }
__device__ __forceinline__ void sts128(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) { asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory"); }
// mode 0: sts64 lo+hi with the kernel's row_of mapping   1: sts64, quarters on rows 0,1,2,3   2: sts64, rows 0,2,4,6
// mode 3: sts128 (16-bit-token layout), kernel mapping    4: sts64 linear (lane*8, conflict-free reference)
__global__ void k(int mode, int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, q = lane >> 3, ql = lane & 7;
  int rows[4];
  for (int st = 0; st < 4; ++st) {
    int r;
    if (mode == 1) r = q + 4 * st; else if (mode == 2) r = (2 * q + (st & 1)) + 8 * (st >> 1);
    else r = ((q & 1) << 2) | (q >> 1) | ((st & 1) << 1) | ((st >> 1) << 3);
    rows[st] = warp * 16 + r;
  }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int st = 0; st < 4; ++st) {
      const uint32_t r = rows[st], sw = r & 7u, c0 = ql >> 1, hb = (ql & 1) * 8;
      if (mode == 3) sts128(base + r * 128 + ((ql ^ sw) << 4), it, st, lane, 0);
      else if (mode == 4) { sts64(base + warp * 2048 + st * 512 + lane * 8, it, st); sts64(base + warp * 2048 + st * 512 + 256 + lane * 8, it, st); }
      else { sts64(base + r * 128 + ((c0 ^ sw) << 4) + hb, it, st); sts64(base + r * 128 + (((c0 + 4) ^ sw) << 4) + hb, it, lane); }
    }
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) out[0] = t1 - t0;
}
int main() {
  long long* d; cudaMalloc(&d, 8);
  for (int warps : {1, 8})
    for (int mode = 0; mode < 5; ++mode) {
      const int iters = 2000;
      k<<<1, 32 * warps, 16384>>>(mode, 10, d);
      k<<<1, 32 * warps, 16384>>>(mode, iters, d);
      long long c; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
      const double bytes = (double)iters * 4 * 256 * 2 * warps / (mode == 3 ? 1 : 1);   // mode 3 stores 512 B per instr, 4 instr
      printf("warps=%d mode=%d  %.2f cycles per 128 B stored (whole SM)\n", warps, mode, (double)c / (bytes / 128.0));
    }
  return 0;
}
