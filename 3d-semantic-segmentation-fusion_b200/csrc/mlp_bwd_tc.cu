// a6 backward on the 5th-generation tensor cores (tcgen05, bf16 operands, fp32 accumulate in TMEM): the gradients of
// the 448->256->64->16 MLP from the bf16 rows / activations the tcgen05 forward saved.  Replaces the autograd of
// FeatureFusionModel.mlp (utils/fusion_model.py:144-150,203-204; utils/train.py:91 calls backward on the total loss).
// Every product is a kernel of this file; no vendor GEMM library is involved.
//
//   dl   = dlogits                               db3 += sum_rows dl          dW3 += dl^T  @ h2
//   dh2  = (dl  @ W3) * (h2 > 0)     [N,64]      db2 += sum_rows dh2         dW2 += dh2^T @ h1
//   dh1  = (dh2 @ W2) * (h1 > 0)     [N,256]     db1 += sum_rows dh1         dW1 += dh1^T @ [x_pt | x_img]
//   dx_pt = dh1 @ W1[:, :64]         [N,64]
//
// Three launches after the operand packing:
//   dgrad_chain_kernel   the data-gradient chain per 128-point tile: dl @ W3, dh2 @ W2 and
//                        dh1 @ W1[:, :64] as tcgen05.mma with the weights resident in shared memory, the ReLU masks in
//                        the TMEM epilogues (the h1 tile arrives by TMA and is overwritten in place by dh1, which is the
//                        A operand of the next product), bias sums and the bf16 copies of dh2 / dh1 for the weight
//                        gradients written from shared memory by two helper warps.  Bound: HBM (about 223 MB per 4 sweeps).
//   dw3_db3_kernel       dW3 (16 x 64, 0.3 GFLOP) and db3 on the CUDA cores, as per-block partials.
//   wgrad_kernel         dW1 and dW2 with the points as the K dimension: both operands are the row-major [points, features]
//                        matrices as they lie in HBM, i.e. MN-major UMMA operands, delivered by TMA into 128B-swizzled
//                        slabs; every CTA accumulates its share of the points into TMEM (512 columns, all used) and writes
//                        one fp32 partial; wgrad_reduce_kernel adds the partials into the caller's buffers in a fixed
//                        order (deterministic).  Two CTA flavours split the 448 input columns so that each accumulator
//                        set fits TMEM:  A: dW1[:, 0:256]   B: dW1[:, 256:448] and dW2^T.
#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>

#include "common.cuh"
#include "tc_ptx.cuh"
#include "tmap.cuh"

namespace tdssf {
namespace bwd {

constexpr int kCp = 64, kCi = 384, kK1 = 448, kN1 = 256, kN2 = 64, kC = 16;

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ void wait_bar(uint32_t bar, uint32_t parity) {
  if (ptx::mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!ptx::mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) __trap();  // ~2 s: a protocol bug, never a legitimate wait
  }
}

// X[r, 0:64] = bf16(x_pt[r]), X[r, 64:448] = bf16(x_img[r]); one thread per 8 columns (only for the unfused forward,
// which saves no bf16 input rows)
__global__ void concat_bf16_kernel(const float* __restrict__ x_pt, const float* __restrict__ x_img, uint4* __restrict__ X, int64_t N) {
  const int64_t total = N * (kK1 / 8);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / (kK1 / 8);
    const int c8 = (int)(i % (kK1 / 8));
    const float* s = c8 < kCp / 8 ? x_pt + r * kCp + c8 * 8 : x_img + r * kCi + (c8 - kCp / 8) * 8;
    const float4 a = __ldg(reinterpret_cast<const float4*>(s)), b = __ldg(reinterpret_cast<const float4*>(s) + 1);
    X[i] = make_uint4(pack2(a.x, a.y), pack2(a.z, a.w), pack2(b.x, b.y), pack2(b.z, b.w));
  }
}

// ---- operand image of the weights for the chain kernel ------------------------------------------------------------
// K-major, 128B-swizzled B operands (row n, element k at n*128 + (((k%64)/8) ^ (n&7))*16 + (k%8)*2 inside a 64-wide K slab):
//   W2t   [256 rows n = layer-1 unit][K = 64 layer-2 units]      value W2[k][n]          32 KB
//   W1pt  4 K slabs of [64 rows n = point-feature column][64 k]  value W1[64*slab+k][n]  4 x 8 KB
//   W3t   [64 rows n = layer-2 unit][K = 16 classes], no swizzle: 8-row x 16-byte core matrices, element (n, k) at
//         (n/8)*256 + (k/8)*128 + (n%8)*16 + (k%8)*2, value W3[k][n]                      2 KB
constexpr int kOffW2t = 0, kOffW1pt = 32768, kOffW3 = 65536, kChainWeightBytes = 65536 + 2048;

__global__ void pack_chain_weights_kernel(const float* __restrict__ W1, const float* __restrict__ W2, const float* __restrict__ W3,
                                          uint8_t* __restrict__ out) {
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < 32768 + 1024; e += gridDim.x * blockDim.x) {
    if (e < 16384) {            // W2t: 256 x 64
      const int n = e >> 6, k = e & 63;
      reinterpret_cast<__nv_bfloat16*>(out + kOffW2t + n * 128 + (((k >> 3) ^ (n & 7)) << 4))[k & 7] = __float2bfloat16_rn(W2[k * kN1 + n]);
    } else if (e < 32768) {     // W1pt: 4 slabs x 64 x 64
      const int i = e - 16384, slab = i >> 12, n = (i >> 6) & 63, k = i & 63;
      reinterpret_cast<__nv_bfloat16*>(out + kOffW1pt + slab * 8192 + n * 128 + (((k >> 3) ^ (n & 7)) << 4))[k & 7] =
          __float2bfloat16_rn(W1[(slab * 64 + k) * kK1 + n]);
    } else {                    // W3t: 64 x 16
      const int i = e - 32768, n = i >> 4, k = i & 15;
      reinterpret_cast<__nv_bfloat16*>(out + kOffW3 + (n >> 3) * 256 + (k >> 3) * 128 + (n & 7) * 16)[k & 7] = __float2bfloat16_rn(W3[k * kN2 + n]);
    }
  }
}

// ---- dW3 = dl^T @ h2 and db3 = column sums of dl on the CUDA cores (2 k FLOP per point: a streaming reduction) ------
// A warp takes 32 consecutive rows at a time: lane l loads dl row l (64 bytes) and, for every row of the group, the h2
// columns 2l, 2l+1 (32 coalesced 128-byte requests), all issued before the first use; the 16 dl values of row j then
// reach every lane by shuffles, and lane l keeps the 16 x 2 sums of its columns in registers.  Per block the eight
// warps are added through shared memory into one partial [1024 + 16]; wgrad_reduce_kernel adds the partials in block
// order (deterministic, no atomics).
constexpr int kW3Out = kC * kN2 + kC;   // 1040
__global__ void __launch_bounds__(256, 2)
dw3_db3_kernel(const float* __restrict__ dl, const uint16_t* __restrict__ h2, int64_t N, float* __restrict__ partial) {
  __shared__ float s_acc[8][kW3Out];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc[kC][2], bs[kC];
#pragma unroll
  for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = bs[c] = 0.f;
  const int64_t n_groups = (N + 31) / 32, W = (int64_t)gridDim.x * 8;
  for (int64_t g = (int64_t)blockIdx.x * 8 + warp; g < n_groups; g += W) {
    const int64_t r0 = g * 32;
    const bool mine = r0 + lane < N;
    const float4* dp = reinterpret_cast<const float4*>(dl + (mine ? r0 + lane : 0) * kC);
    float4 d4[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) d4[k] = mine ? __ldg(dp + k) : make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t hv[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) hv[j] = r0 + j < N ? __ldg(reinterpret_cast<const uint32_t*>(h2 + (r0 + j) * kN2) + lane) : 0u;
    const float d[kC] = {d4[0].x, d4[0].y, d4[0].z, d4[0].w, d4[1].x, d4[1].y, d4[1].z, d4[1].w,
                         d4[2].x, d4[2].y, d4[2].z, d4[2].w, d4[3].x, d4[3].y, d4[3].z, d4[3].w};
#pragma unroll
    for (int c = 0; c < kC; ++c) bs[c] += d[c];   // this lane's row; the lanes are added at the end
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float h0 = bf_lo(hv[j]), h1 = bf_hi(hv[j]);
#pragma unroll
      for (int c = 0; c < kC; ++c) {
        const float v = __shfl_sync(0xffffffffu, d[c], j);
        acc[c][0] = fmaf(v, h0, acc[c][0]);
        acc[c][1] = fmaf(v, h1, acc[c][1]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < kC; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bs[c] += __shfl_xor_sync(0xffffffffu, bs[c], o);
    *reinterpret_cast<float2*>(&s_acc[warp][c * kN2 + 2 * lane]) = make_float2(acc[c][0], acc[c][1]);
    if (lane == 0) s_acc[warp][kC * kN2 + c] = bs[c];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < kW3Out; e += 256) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += s_acc[w][e];
    partial[(size_t)blockIdx.x * kW3Out + e] = v;
  }
}

// =====================================================================================================================
// dgrad_chain_kernel: per 128-point tile   dh2 = (dl @ W3) * (h2 > 0)   dh1 = (dh2 @ W2) * (h1 > 0)   dx_pt = dh1 @ W1[:, :64]
// One CTA per SM (cta_group::1), persistent over tiles, 14 working warps:
//   warps 0-3   epilogue, one tile row per thread (TMEM lane): acc1 -> ReLU mask from the h1 tile in shared memory -> bf16
//               dh1 written over h1 in place (the 128B-swizzled K-major slabs are the A operand of the next product);
//               accx -> dx_pt (fp32) to HBM
//   warps 4-7   dh2 of the NEXT tile, one row per thread: dl -> bf16 -> K = 16 A operand (no swizzle) of dl @ W3 on the tensor
//               core; then its epilogue: mask from h2, bf16 into the dh2 slab; the rows' dl and h2 are fetched one tile ahead
//   warp  8     TMEM allocation and tcgen05.mma issue: dl @ W3 (N = 64, K = 16), dh2 @ W2 (N = 256, K = 64), dh1 @ W1[:, :64]
//               (N = 64, K = 256)
//   warp  9     TMA: the tile's h1 rows as four [128 x 64] slabs (zero fill beyond N), two tiles in flight; every slab has
//               its own full / free barriers, so it is refilled as soon as its K step and its copy-out are done
//   warps 10-13 copy dh1 (one slab each) and dh2 (warp 10) from the slabs to HBM with coalesced stores (the weight-
//               gradient kernel reads them) and accumulate the bias gradients db1 / db2 (column sums) in registers
// =====================================================================================================================
namespace chain {
#ifdef TDSSF_CH_TRACE   // (experiment builds) clock64 of event `ev` of tile `it`, CTA 0 only
#define CH_TRACE(ev, it)                                                                                   \
  do {                                                                                                     \
    if (p.trace && blockIdx.x == 0 && (threadIdx.x & 31) == 0 && (it) < 16) p.trace[(ev) * 16 + (it)] = clock64(); \
  } while (0)
#else
#define CH_TRACE(ev, it)
#endif
constexpr int kThreads = 512;
constexpr int kOffH1 = kChainWeightBytes;           // 69632, 1024-aligned; two buffers of 4 x 16 KB
constexpr int kOffDh2 = kOffH1 + 2 * 65536;         // 198656
constexpr int kOffDl = kOffDh2 + 16384;             // dl tile as a no-swizzle K-major A operand: 128 rows x 32 bytes = 4 KB
constexpr int kOffBar = kOffDl + 4096;
constexpr int kOffTmem = kOffBar + 32 * 8;
constexpr int kSmemBytes = kOffTmem + 16;
static_assert(kOffH1 % 1024 == 0 && kOffDh2 % 1024 == 0, "slab alignment");
static_assert(kSmemBytes + 1024 <= 232448, "shared memory budget");
// the h1 / dh1 tile is handed over slab by slab (64 columns): [buffer][slab] full / free barriers, one dh1-full barrier per slab
enum { kH1Full = 0, kH1Free = 8, kDh2Full = 16, kDh2Free = 17, kAcc1Full = 18, kAcc1Free = 19, kDh1Full = 20, kAccxFull = 24, kAccxFree = 25,
       kWFull = 26, kDlFull = 27, kAcc2Full = 28, kNumBars = 29 };

struct Params {
  const float* dl;          // [N,16] fp32
  const uint16_t* h2;       // [N,64] bf16
  const uint8_t* weights;   // pack_chain_weights_kernel image
  uint16_t* dh2_out;        // [N,64] bf16
  uint16_t* dh1_out;        // [N,256] bf16
  float* dx_pt;             // [N,64] fp32 or NULL
  float* bias_partial;      // [grid][320]: this CTA's column sums of dh1 (256) and dh2 (64); summed by wgrad_reduce_kernel
  int64_t N;
  int n_tiles;
  long long* trace;         // experiment builds only
};

// 0xffff per half of a packed bf16 pair that is > 0 (the ReLU mask of the saved activation)
__device__ __forceinline__ uint32_t relu_mask2(uint32_t h) {
  return __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&h), __floats2bfloat162_rn(0.f, 0.f));
}

__global__ void __launch_bounds__(kThreads, 1)
dgrad_chain_kernel(const __grid_constant__ CUtensorMap tm_h1, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto bar = [&](int i) { return sbase + kOffBar + 8 * i; };
  const int n_iters = p.n_tiles > (int)blockIdx.x ? (p.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  auto tile_of = [&](int it) { return (int64_t)blockIdx.x + (int64_t)it * gridDim.x; };

  if (threadIdx.x == 0) {
    for (int b = 0; b < 8; ++b) { ptx::mbar_init(bar(kH1Full + b), 1); ptx::mbar_init(bar(kH1Free + b), 2); }   // commit + copy-out warp
    ptx::mbar_init(bar(kDh2Full), 4);
    ptx::mbar_init(bar(kDh2Free), 5);   // the product's commit + the four copy-out warps
    ptx::mbar_init(bar(kWFull), 1);
    ptx::mbar_init(bar(kDlFull), 4);
    ptx::mbar_init(bar(kAcc2Full), 1);
    ptx::mbar_init(bar(kAcc1Full), 1);
    ptx::mbar_init(bar(kAcc1Free), 4);
    for (int b = 0; b < 4; ++b) ptx::mbar_init(bar(kDh1Full + b), 4);
    ptx::mbar_init(bar(kAccxFull), 1);
    ptx::mbar_init(bar(kAccxFree), 4);
    ptx::fence_mbar_init();
    ptx::tma_prefetch_desc(&tm_h1);
  }
  if (warp == 8) {
    ptx::tmem_alloc_1cta(sbase + kOffTmem, 512);
    ptx::tmem_relinquish_1cta();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem + kOffTmem);
  if (threadIdx.x == 0) {   // the weight image lands (TMA bulk copies) while the first tile's inputs are fetched
    ptx::mbar_expect_tx(bar(kWFull), kChainWeightBytes);
    for (int off = 0; off < kChainWeightBytes; off += 16384)
      ptx::bulk_g2s(sbase + off, p.weights + off, min(16384, kChainWeightBytes - off), bar(kWFull));
  }

  if (warp < 4) {
    // ------------------------------------------------ epilogue -------------------------------------------------------
    const int r = threadIdx.x;
    const uint32_t tl = tmem + (((uint32_t)warp * 32u) << 16);
    const uint32_t sw = (uint32_t)r & 7u;
    for (int it = 0; it < n_iters; ++it) {
      const int hb = it & 1;
      const int64_t row = tile_of(it) * 128 + r;
      wait_bar(bar(kAcc1Full), it & 1);
      ptx::tc_fence_after();
      if (warp == 0) CH_TRACE(0, it);
      uint32_t va[16], vb[16];
      uint4 hq[8];   // the row's eight 16-byte chunks of the current h1 slab, loaded in one go when the slab has landed
      ptx::tmem_ld16(tl, va);
#pragma unroll
      for (int c = 0; c < 16; c += 2) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t(&v)[16] = half ? vb : va;
          const int cc = c + half;
          const uint32_t slab = sbase + kOffH1 + (uint32_t)hb * 65536u + (uint32_t)(cc >> 2) * 16384u + (uint32_t)r * 128u;
          if ((cc & 3) == 0) {
            wait_bar(bar(kH1Full + hb * 4 + (cc >> 2)), (it >> 1) & 1);   // this slab of h1 has landed
#pragma unroll
            for (int q = 0; q < 8; ++q) hq[q] = ptx::lds128(slab + (((uint32_t)q ^ sw) << 4));
          }
          ptx::tmem_wait_ld();
          if (half == 0) ptx::tmem_ld16(tl + (c + 1) * 16, vb);
          else if (c + 2 < 16) ptx::tmem_ld16(tl + (c + 2) * 16, va);
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const int q = (cc & 3) * 2 + k;
            const uint4 h = hq[q];
            ptx::sts128(slab + (((uint32_t)q ^ sw) << 4), pack2(__uint_as_float(v[8 * k]), __uint_as_float(v[8 * k + 1])) & relu_mask2(h.x),
                        pack2(__uint_as_float(v[8 * k + 2]), __uint_as_float(v[8 * k + 3])) & relu_mask2(h.y),
                        pack2(__uint_as_float(v[8 * k + 4]), __uint_as_float(v[8 * k + 5])) & relu_mask2(h.z),
                        pack2(__uint_as_float(v[8 * k + 6]), __uint_as_float(v[8 * k + 7])) & relu_mask2(h.w));
          }
          if ((cc & 3) == 3) {   // a dh1 slab is complete: the next product's K step and the copy-out warp may read it
            if (cc == 15) ptx::tc_fence_before();
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (cc == 15) ptx::mbar_arrive_local(bar(kAcc1Free));
              ptx::mbar_arrive_local(bar(kDh1Full + (cc >> 2)));
            }
            __syncwarp();
          }
        }
      }
      if (warp == 0) CH_TRACE(1, it);
      // dx_pt
      wait_bar(bar(kAccxFull), it & 1);
      ptx::tc_fence_after();
      if (warp == 0) CH_TRACE(2, it);
      uint32_t x0[16], x1[16];
      uint4* dst = reinterpret_cast<uint4*>(p.dx_pt + (row < p.N ? row : 0) * kCp);
      const bool store = p.dx_pt != nullptr && row < p.N;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        ptx::tmem_ld16(tl + 256 + half * 32, x0);
        ptx::tmem_ld16(tl + 272 + half * 32, x1);
        ptx::tmem_wait_ld();
        if (half == 1) {   // the accumulator is in registers: the next tile's product may overwrite it
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive_local(bar(kAccxFree));
          __syncwarp();
        }
        if (store) {   // a row per thread: 32-byte stores, every lane in its own line
          ptx::stg256(dst + half * 8, x0[0], x0[1], x0[2], x0[3], x0[4], x0[5], x0[6], x0[7]);
          ptx::stg256(dst + half * 8 + 2, x0[8], x0[9], x0[10], x0[11], x0[12], x0[13], x0[14], x0[15]);
          ptx::stg256(dst + half * 8 + 4, x1[0], x1[1], x1[2], x1[3], x1[4], x1[5], x1[6], x1[7]);
          ptx::stg256(dst + half * 8 + 6, x1[8], x1[9], x1[10], x1[11], x1[12], x1[13], x1[14], x1[15]);
        }
      }
      if (warp == 0) CH_TRACE(3, it);
    }
  } else if (warp < 8) {
    // ------------------------------------------------ dh2 producers --------------------------------------------------
    const int i = threadIdx.x - 128;
    const uint32_t sw = (uint32_t)i & 7u;
    const uint32_t slab = sbase + kOffDh2 + (uint32_t)i * 128u;
    // dl (fp32) -> bf16 -> the K = 16 A operand of dl @ W3 (one tcgen05.mma per tile, accumulator columns [320,384));
    // then this warp is the epilogue of that product for its 32 rows: mask from h2, bf16, into the dh2 slab.  The next
    // tile's dl / h2 rows are fetched while the product and the mask run.
    ptx::u32x8 dn[2], hn[4];
    auto prefetch = [&](int it) {
      const int64_t row = tile_of(it) * 128 + i;
      const bool live = it < n_iters && row < p.N;
      const char* g = reinterpret_cast<const char*>(p.dl + (live ? row : 0) * kC);
      const char* hh = reinterpret_cast<const char*>(p.h2 + (live ? row : 0) * kN2);
#pragma unroll
      for (int k = 0; k < 2; ++k) dn[k] = ptx::ldg256(g + 32 * k);
#pragma unroll
      for (int k = 0; k < 4; ++k) hn[k] = ptx::ldg256(hh + 32 * k);
      if (!live) {   // rows beyond N: zero gradient and zero mask
#pragma unroll
        for (int e = 0; e < 8; ++e) { dn[0].v[e] = dn[1].v[e] = 0u; hn[0].v[e] = hn[1].v[e] = hn[2].v[e] = hn[3].v[e] = 0u; }
      }
    };
    const uint32_t tl = tmem + (((uint32_t)(warp & 3) * 32u) << 16) + 320u;
    const uint32_t dl_row = sbase + kOffDl + (uint32_t)(i >> 3) * 256u + (uint32_t)(i & 7) * 16u;
    prefetch(0);
    for (int it = 0; it < n_iters; ++it) {
      if (warp == 4) CH_TRACE(4, it);
      // (the previous tile's product has completed: this warp waited for its accumulator, so the dl operand is free)
#pragma unroll
      for (int k = 0; k < 2; ++k)
        ptx::sts128(dl_row + 128u * k, pack2(__uint_as_float(dn[k].v[0]), __uint_as_float(dn[k].v[1])),
                    pack2(__uint_as_float(dn[k].v[2]), __uint_as_float(dn[k].v[3])), pack2(__uint_as_float(dn[k].v[4]), __uint_as_float(dn[k].v[5])),
                    pack2(__uint_as_float(dn[k].v[6]), __uint_as_float(dn[k].v[7])));
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_local(bar(kDlFull));
      __syncwarp();
      uint32_t hm[32];   // the row's 64 ReLU masks (packed pairs); frees hn for the next tile's loads
#pragma unroll
      for (int k = 0; k < 32; ++k) hm[k] = relu_mask2(hn[k >> 3].v[k & 7]);
      prefetch(it + 1);
      wait_bar(bar(kAcc2Full), it & 1);
      ptx::tc_fence_after();
      if (warp == 4) CH_TRACE(5, it);
      uint32_t res[32];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t v[16];
        ptx::tmem_ld16(tl + c * 16, v);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 8; ++k) res[8 * c + k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1])) & hm[8 * c + k];
      }
      ptx::tc_fence_before();
      if (it >= 1) wait_bar(bar(kDh2Free), (it - 1) & 1);
      if (warp == 4) CH_TRACE(6, it);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        ptx::sts128(slab + (((uint32_t)q ^ sw) << 4), res[4 * q], res[4 * q + 1], res[4 * q + 2], res[4 * q + 3]);
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_local(bar(kDh2Full));
      __syncwarp();
      if (warp == 4) CH_TRACE(7, it);
    }
  } else if (warp == 8) {
    // ------------------------------------------------ MMA issue ------------------------------------------------------
    if (lane == 0) {
      constexpr uint32_t kI1 = ptx::idesc_bf16(128, kN1), kIx = ptx::idesc_bf16(128, kCp), kI2 = ptx::idesc_bf16(128, kN2);
      // dl @ W3 of tile `it`: one K = 16 instruction on the no-swizzle operands, accumulator columns [320,384)
      auto issue_l2 = [&](int it) {
        wait_bar(bar(kDlFull), it & 1);
        ptx::tc_fence_after();
        ptx::umma_ss_1cta(tmem + 320, ptx::smem_desc_nosw(sbase + kOffDl, 128, 256), ptx::smem_desc_nosw(sbase + kOffW3, 128, 256), kI2, 0);
        ptx::umma_commit_1cta(bar(kAcc2Full));
      };
      wait_bar(bar(kWFull), 0);
      if (n_iters > 0) issue_l2(0);
      for (int it = 0; it < n_iters; ++it) {
        const int hb = it & 1;
        wait_bar(bar(kDh2Full), it & 1);
        // the next tile's dl @ W3 first: its operand arrives right after this tile's dh2 slab, and its accumulator was read
        // before that slab was announced - the producers' masks then overlap the wait for the previous tile's epilogue
        if (it + 1 < n_iters) issue_l2(it + 1);
        if (it >= 1) wait_bar(bar(kAcc1Free), (it - 1) & 1);
        ptx::tc_fence_after();
        CH_TRACE(8, it);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          ptx::umma_ss_1cta(tmem, ptx::smem_desc_sw128(sbase + kOffDh2 + kk * 32), ptx::smem_desc_sw128(sbase + kOffW2t + kk * 32), kI1, kk != 0);
        ptx::umma_commit_1cta(bar(kAcc1Full));
        ptx::umma_commit_1cta(bar(kDh2Free));
        CH_TRACE(9, it);
        if (it >= 1) wait_bar(bar(kAccxFree), (it - 1) & 1);
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          wait_bar(bar(kDh1Full + s), it & 1);
          ptx::tc_fence_after();
          if (s == 0) CH_TRACE(10, it);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            ptx::umma_ss_1cta(tmem + 256, ptx::smem_desc_sw128(sbase + kOffH1 + hb * 65536 + s * 16384 + kk * 32),
                              ptx::smem_desc_sw128(sbase + kOffW1pt + s * 8192 + kk * 32), kIx, (s | kk) != 0);
          ptx::umma_commit_1cta(bar(kH1Free + hb * 4 + s));
        }
        ptx::umma_commit_1cta(bar(kAccxFull));
        CH_TRACE(11, it);
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ------------------------------------------------ TMA: h1 tiles --------------------------------------------------
    if (lane == 0) {
      for (int it = 0; it < n_iters; ++it) {
        const int hb = it & 1;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          if (it >= 2) wait_bar(bar(kH1Free + hb * 4 + s), ((it >> 1) - 1) & 1);
          if (s == 0) CH_TRACE(16, it);
          ptx::mbar_expect_tx(bar(kH1Full + hb * 4 + s), 16384);
          ptx::tma_load_2d(sbase + kOffH1 + hb * 65536 + s * 16384, &tm_h1, 64 * s, (int)(tile_of(it) * 128), bar(kH1Full + hb * 4 + s));
        }
      }
    }
    __syncwarp();
  } else if (warp < 14) {
    // ------------------------------------------------ copy-out + bias sums -------------------------------------------
    const int h = warp - 10, q = lane & 7, rg = lane >> 3;
    float s1[8], s2[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) s1[e] = s2[e] = 0.f;
    // rows [r_lo, r_hi) of a [128 x 64] slab: four rows per instruction (eight lanes x 16 bytes each)
    auto drain = [&](uint32_t slab, uint16_t* out, int ld, int col0, int64_t row0, int r_lo, int r_hi, float (&acc)[8]) {
      for (int r4 = r_lo + rg; r4 < r_hi; r4 += 16) {   // four loads in flight, then their sums and stores
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int r = r4 + 4 * u;
          v[u] = ptx::lds128(slab + (uint32_t)r * 128u + ((((uint32_t)q) ^ ((uint32_t)r & 7u)) << 4));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int r = r4 + 4 * u;
          acc[0] += bf_lo(v[u].x); acc[1] += bf_hi(v[u].x); acc[2] += bf_lo(v[u].y); acc[3] += bf_hi(v[u].y);
          acc[4] += bf_lo(v[u].z); acc[5] += bf_hi(v[u].z); acc[6] += bf_lo(v[u].w); acc[7] += bf_hi(v[u].w);
          if (row0 + r < p.N) *reinterpret_cast<uint4*>(out + (row0 + r) * ld + col0 + q * 8) = v[u];
        }
      }
    };
    for (int it = 0; it < n_iters; ++it) {
      const int hb = it & 1;
      const int64_t row0 = tile_of(it) * 128;
      wait_bar(bar(kDh2Full), it & 1);
      if (h == 0) CH_TRACE(12, it);
      drain(sbase + kOffDh2, p.dh2_out, kN2, 0, row0, 32 * h, 32 * h + 32, s2);
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_local(bar(kDh2Free));
      __syncwarp();
      if (h == 0) CH_TRACE(13, it);
      wait_bar(bar(kDh1Full + h), it & 1);
      if (h == 0) CH_TRACE(14, it);
      drain(sbase + kOffH1 + hb * 65536 + h * 16384, p.dh1_out, kN1, h * 64, row0, 0, 128, s1);
      if (h == 0) CH_TRACE(15, it);
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_local(bar(kH1Free + hb * 4 + h));
      __syncwarp();
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      s1[e] += __shfl_xor_sync(0xffffffffu, s1[e], 8);
      s1[e] += __shfl_xor_sync(0xffffffffu, s1[e], 16);
      s2[e] += __shfl_xor_sync(0xffffffffu, s2[e], 8);
      s2[e] += __shfl_xor_sync(0xffffffffu, s2[e], 16);
    }
    // per-CTA partial sums (no atomics: 148 CTAs finishing together would serialise on 320 addresses); db2 of the four
    // copy-out warps goes through shared memory (the dl operand buffer is idle by now)
    float* s_db2 = reinterpret_cast<float*>(smem + kOffDl);
    float* bp = p.bias_partial + (size_t)blockIdx.x * 320;
    if (rg == 0) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        bp[h * 64 + q * 8 + e] = s1[e];
        s_db2[h * 64 + q * 8 + e] = s2[e];
      }
    }
    ptx::named_bar_sync(1, 128);
    if (h == 0) {
      for (int c = lane; c < 64; c += 32) bp[256 + c] = (s_db2[c] + s_db2[64 + c]) + (s_db2[128 + c] + s_db2[192 + c]);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 8) ptx::tmem_dealloc_1cta(tmem, 512);
}
}  // namespace chain

// =====================================================================================================================
// wgrad_kernel: dW1 = dh1^T @ X and dW2^T = h1^T @ dh2 with the points as the K dimension.
// A stage is 64 points: every operand slab is [64 points x 64 features] bf16 = 8 KB, delivered by one TMA box load into
// the 128B-swizzled layout, which for a [k = point][m/n = feature] tile IS the canonical MN-major UMMA atom
// (cute/atom/mma_traits_sm100.hpp: Swizzle<3,4,3> o ((T,8,m),(8,k)):((1,T,LBO),(8T,SBO)) with LBO = slab size, SBO = 1024).
// Rows beyond N are zero-filled by TMA and contribute nothing.
//   flavour A (CTAs [0, n_a)):      slabs dh1 x4, X[:, 0:256] x4           8 slabs / stage, 3 stages
//       TMEM [0,256) = dW1[0:128, 0:256]      [256,512) = dW1[128:256, 0:256]
//   flavour B (CTAs [n_a, grid)):   slabs dh1 x4, X[:, 256:448] x3, h1 x4, dh2 x1     12 slabs / stage, 2 stages
//       TMEM [0,192) = dW1[0:128, 256:448]    [192,384) = dW1[128:256, 256:448]
//            [384,448) = dW2^T[0:128, :]      [448,512) = dW2^T[128:256, :]
// Warp 0 issues the TMA loads, warp 1 the MMAs; at the end warps 2-9 move the accumulators TMEM -> shared memory
// (transposed, 32 x 32 blocks) -> the CTA's fp32 partial [128][512] with coalesced 16-byte stores.
// =====================================================================================================================
namespace wg {
constexpr int kKP = 64;                 // points per stage
constexpr int kSlab = kKP * 128;        // 8192
constexpr int kThreads = 320;
constexpr int kRing = 24 * kSlab;       // 196608: 3 stages x 8 slabs or 2 stages x 12 slabs
constexpr int kOffBar = kRing;
constexpr int kOffTmem = kOffBar + 16 * 8;
constexpr int kSmemBytes = kOffTmem + 16;
constexpr int kPartialFloats = 128 * 512;

struct Params {
  float* partials;     // [grid][128][512]
  int64_t N;
  int n_stages;        // ceil(N / 64)
  int n_a;             // CTAs of flavour A
};

template <bool kB>
__device__ __forceinline__ void run(const CUtensorMap& tm_dh1, const CUtensorMap& tm_x, const CUtensorMap& tm_h1, const CUtensorMap& tm_dh2,
                                    const Params& p, uint8_t* smem, uint32_t sbase, uint32_t tmem, int warp, int lane) {
  constexpr int kStages = kB ? 2 : 3, kSlabs = kB ? 12 : 8;
  constexpr uint32_t kStageBytes = kSlabs * kSlab;
  auto bar = [&](int i) { return sbase + kOffBar + 8 * i; };   // [0,kStages) full, [4,4+kStages) empty, 8 done
  const int first = kB ? (int)blockIdx.x - p.n_a : (int)blockIdx.x;
  const int stride = kB ? (int)gridDim.x - p.n_a : p.n_a;
  const int n_my = p.n_stages > first ? (p.n_stages - first + stride - 1) / stride : 0;

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < n_my; ++i) {
        const int st = i % kStages, row0 = (first + i * stride) * kKP;
        if (i >= kStages) wait_bar(bar(4 + st), ((i / kStages) - 1) & 1);
        ptx::mbar_expect_tx(bar(st), kStageBytes);
        const uint32_t base = sbase + st * kStageBytes;
#pragma unroll
        for (int s = 0; s < 4; ++s) ptx::tma_load_2d(base + s * kSlab, &tm_dh1, 64 * s, row0, bar(st));
        if (!kB) {
#pragma unroll
          for (int s = 0; s < 4; ++s) ptx::tma_load_2d(base + (4 + s) * kSlab, &tm_x, 64 * s, row0, bar(st));
        } else {
#pragma unroll
          for (int s = 0; s < 3; ++s) ptx::tma_load_2d(base + (4 + s) * kSlab, &tm_x, 256 + 64 * s, row0, bar(st));
#pragma unroll
          for (int s = 0; s < 4; ++s) ptx::tma_load_2d(base + (7 + s) * kSlab, &tm_h1, 64 * s, row0, bar(st));
          ptx::tma_load_2d(base + 11 * kSlab, &tm_dh2, 0, row0, bar(st));
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t kMN = ptx::kIdescAMajorMN | ptx::kIdescBMajorMN;
      constexpr uint32_t kI256 = ptx::idesc_bf16(128, 256) | kMN, kI192 = ptx::idesc_bf16(128, 192) | kMN, kI64 = ptx::idesc_bf16(128, 64) | kMN;
      for (int i = 0; i < n_my; ++i) {
        const int st = i % kStages;
        wait_bar(bar(st), (i / kStages) & 1);
        ptx::tc_fence_after();
        const uint32_t base = sbase + st * kStageBytes;
#pragma unroll
        for (int kk = 0; kk < kKP / 16; ++kk) {
          const uint32_t koff = kk * 2048;   // 16 k-rows of 128 bytes
          const uint32_t acc = (i | kk) != 0;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint64_t a = ptx::smem_desc_sw128_ex(base + (2 * h) * kSlab + koff, kSlab, 1024);
            if (!kB) {
              ptx::umma_ss_1cta(tmem + h * 256, a, ptx::smem_desc_sw128_ex(base + 4 * kSlab + koff, kSlab, 1024), kI256, acc);
            } else {
              ptx::umma_ss_1cta(tmem + h * 192, a, ptx::smem_desc_sw128_ex(base + 4 * kSlab + koff, kSlab, 1024), kI192, acc);
              ptx::umma_ss_1cta(tmem + 384 + h * 64, ptx::smem_desc_sw128_ex(base + (7 + 2 * h) * kSlab + koff, kSlab, 1024),
                                ptx::smem_desc_sw128_ex(base + 11 * kSlab + koff, kSlab, 1024), kI64, acc);
            }
          }
        }
        ptx::umma_commit_1cta(bar(4 + st));
      }
      ptx::umma_commit_1cta(bar(8));
    }
    __syncwarp();
  } else {
    // epilogue, eight warps: warp w owns TMEM lanes 32*(w & 3) and the column half (w - 2) >> 2; blocks of 32 columns go
    // through a [32][36]-float transpose buffer (16-byte accesses on both sides, conflict-free: rows 144 bytes apart) so
    // that the partial is written with 16-byte stores that cover four full 128-byte rows per instruction
    wait_bar(bar(8), 0);
    ptx::tc_fence_after();
    const int qd = warp & 3, ch = (warp - 2) >> 2;
    float* buf = reinterpret_cast<float*>(smem) + (warp - 2) * (32 * 36);   // the ring is idle once `done` completed
    float* out = p.partials + (size_t)blockIdx.x * kPartialFloats + (size_t)(qd * 32) * 512;
    const uint32_t tl = tmem + (((uint32_t)qd * 32u) << 16);
    for (int c0 = ch * 256; c0 < ch * 256 + 256; c0 += 32) {
      uint32_t va[16], vb[16];
      if (n_my > 0) {
        ptx::tmem_ld16(tl + c0, va);
        ptx::tmem_ld16(tl + c0 + 16, vb);
        ptx::tmem_wait_ld();
      } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) va[k] = vb[k] = 0u;
      }
      __syncwarp();
      uint4* mine = reinterpret_cast<uint4*>(buf + lane * 36);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        mine[k] = make_uint4(va[4 * k], va[4 * k + 1], va[4 * k + 2], va[4 * k + 3]);
        mine[4 + k] = make_uint4(vb[4 * k], vb[4 * k + 1], vb[4 * k + 2], vb[4 * k + 3]);
      }
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int r = 4 * j + (lane >> 3);
        *reinterpret_cast<uint4*>(out + (size_t)r * 512 + c0 + 4 * (lane & 7)) = *reinterpret_cast<const uint4*>(buf + r * 36 + 4 * (lane & 7));
      }
    }
  }
}

__global__ void __launch_bounds__(kThreads, 1)
wgrad_kernel(const __grid_constant__ CUtensorMap tm_dh1, const __grid_constant__ CUtensorMap tm_x,
             const __grid_constant__ CUtensorMap tm_h1, const __grid_constant__ CUtensorMap tm_dh2, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 9; ++i) ptx::mbar_init(sbase + kOffBar + 8 * i, 1);
    ptx::fence_mbar_init();
    ptx::tma_prefetch_desc(&tm_dh1);
    ptx::tma_prefetch_desc(&tm_x);
    ptx::tma_prefetch_desc(&tm_h1);
    ptx::tma_prefetch_desc(&tm_dh2);
  }
  if (warp == 1) {
    ptx::tmem_alloc_1cta(sbase + kOffTmem, 512);
    ptx::tmem_relinquish_1cta();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem + kOffTmem);
  if ((int)blockIdx.x < p.n_a) run<false>(tm_dh1, tm_x, tm_h1, tm_dh2, p, smem, sbase, tmem, warp, lane);
  else run<true>(tm_dh1, tm_x, tm_h1, tm_dh2, p, smem, sbase, tmem, warp, lane);
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc_1cta(tmem, 512);
}

// dW1[f][x] += sum over the flavour-A partials (x < 256) or flavour-B partials (x >= 256); dW2[g][f] += flavour-B partials.
// One thread per output element, partials added in CTA order (deterministic).
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ partials, int n_a, int n_cta, float* __restrict__ dW1, float* __restrict__ dW2,
                    const float* __restrict__ partials3, int n3, float* __restrict__ dW3, float* __restrict__ db3,
                    const float* __restrict__ bias_partials, int n_chain, float* __restrict__ db1, float* __restrict__ db2) {
  constexpr int kMainBlocks = (kN1 * kK1 + kN1 * kN2 + 255) / 256;
  constexpr int kW3Blocks = (kW3Out + 31) / 32;
  if ((int)blockIdx.x >= kMainBlocks + kW3Blocks) {
    // db1 / db2: 32 outputs per block from the chain kernel's per-CTA column sums, same scheme as dW3 below
    __shared__ float s[8][32];
    const int o = ((int)blockIdx.x - kMainBlocks - kW3Blocks) * 32 + (threadIdx.x & 31), k = threadIdx.x >> 5;
    float v = 0.f;
    for (int b = k; b < n_chain; b += 8) v += bias_partials[(size_t)b * 320 + o];
    s[k][threadIdx.x & 31] = v;
    __syncthreads();
    if (k == 0) {
      float t = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) t += s[j][threadIdx.x];
      if (o < 256) db1[o] += t; else db2[o - 256] += t;
    }
    return;
  }
  if ((int)blockIdx.x >= kMainBlocks) {
    // dW3 / db3: 32 outputs per block, eight threads per output walk the dw3_db3_kernel partials, then a shared-memory sum
    __shared__ float s[8][32];
    const int o = ((int)blockIdx.x - kMainBlocks) * 32 + (threadIdx.x & 31), k = threadIdx.x >> 5;
    float v = 0.f;
    if (o < kW3Out)
      for (int b = k; b < n3; b += 8) v += partials3[(size_t)b * kW3Out + o];
    s[k][threadIdx.x & 31] = v;
    __syncthreads();
    if (k == 0 && o < kW3Out) {
      float t = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) t += s[j][threadIdx.x];
      if (o < kC * kN2) dW3[o] += t; else db3[o - kC * kN2] += t;
    }
    return;
  }
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e < kN1 * kK1) {
    const int f = e / kK1, x = e % kK1;
    const bool a = x < 256;
    const float* src = partials + (size_t)(a ? 0 : n_a) * kPartialFloats + (size_t)(f & 127) * 512 + (a ? (f >> 7) * 256 + x : (f >> 7) * 192 + (x - 256));
    const int n = a ? n_a : n_cta - n_a;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int i = 0;
    for (; i + 4 <= n; i += 4) {
      s0 += src[(size_t)i * kPartialFloats];
      s1 += src[(size_t)(i + 1) * kPartialFloats];
      s2 += src[(size_t)(i + 2) * kPartialFloats];
      s3 += src[(size_t)(i + 3) * kPartialFloats];
    }
    for (; i < n; ++i) s0 += src[(size_t)i * kPartialFloats];
    dW1[e] += (s0 + s1) + (s2 + s3);
  } else if (e < kN1 * kK1 + kN1 * kN2) {
    const int i2 = e - kN1 * kK1, f = i2 / kN2, g = i2 % kN2;
    const float* src = partials + (size_t)n_a * kPartialFloats + (size_t)(f & 127) * 512 + 384 + (f >> 7) * 64 + g;
    float s = 0.f;
    for (int i = 0; i < n_cta - n_a; ++i) s += src[(size_t)i * kPartialFloats];
    dW2[g * kN1 + f] += s;
  }
}
}  // namespace wg

// ---- host side ------------------------------------------------------------------------------------------------------
constexpr int kMaxW3Blocks = 1024;
struct Work {
  uint16_t *X, *dh2, *dh1;
  uint8_t* chain_w;
  float* partials;
  float* partials3;
  float* bias_partials;
};
static size_t al(size_t b) { return (b + 255) & ~(size_t)255; }
static size_t carve(int64_t N, int n_cta, char* base, Work* w) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += al(bytes); return p; };
  uint16_t* X = (uint16_t*)take((size_t)N * kK1 * 2);
  uint16_t* dh2 = (uint16_t*)take((size_t)N * kN2 * 2);
  uint16_t* dh1 = (uint16_t*)take((size_t)N * kN1 * 2);
  uint8_t* cw = (uint8_t*)take(kChainWeightBytes);
  float* partials = (float*)take((size_t)n_cta * wg::kPartialFloats * 4);
  float* partials3 = (float*)take((size_t)kMaxW3Blocks * kW3Out * 4);
  float* bias_partials = (float*)take((size_t)1024 * 320 * 4);
  if (w) *w = Work{X, dh2, dh1, cw, partials, partials3, bias_partials};
  return off;
}
// the weight-gradient grid: one CTA per SM, at most one per stage of 64 points and at least one of each flavour
static int wgrad_ctas(int64_t N) { return (int)std::max<int64_t>(2, std::min<int64_t>(num_sms(), 2 * ceil_div(N, wg::kKP))); }

template <typename K>
static int set_smem(K kern, int bytes, bool (&done)[kMaxDevices]) {
  const int dev = current_device();
  if (!done[dev]) {
    TDSSF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    done[dev] = true;
  }
  return TDSSF_OK;
}

}  // namespace bwd
}  // namespace tdssf

extern "C" size_t tdssf_mlp_bwd_tc_workspace_bytes(int64_t N) {
  if (N < 0) N = 0;
  return tdssf::bwd::carve(N, tdssf::bwd::wgrad_ctas(N), nullptr, nullptr) + 256;
}

extern "C" int tdssf_mlp_bwd_tc(const float* x_pt, const float* x_img, const void* x_bf16, const void* h1_bf16, const void* h2_bf16,
                                const float* dlogits, int64_t N, const tdssf_mlp_weights* w_host,
                                const tdssf_mlp_grads* g_host, float* dx_pt, void* workspace, size_t workspace_bytes,
                                void* stream) {
  using namespace tdssf;
  using namespace tdssf::bwd;
  TDSSF_REQUIRE((x_bf16 || (x_pt && x_img)) && h1_bf16 && h2_bf16 && dlogits && w_host && g_host, "tdssf_mlp_bwd_tc: NULL pointer");
  TDSSF_REQUIRE(w_host->W1 && w_host->W2 && w_host->W3, "tdssf_mlp_bwd_tc: NULL weight pointer");
  TDSSF_REQUIRE(g_host->dW1 && g_host->db1 && g_host->dW2 && g_host->db2 && g_host->dW3 && g_host->db3,
                "tdssf_mlp_bwd_tc: NULL gradient pointer");
  TDSSF_REQUIRE(N >= 0 && N < (1ll << 31) - 256, "tdssf_mlp_bwd_tc: N out of range");
  if (N == 0) return TDSSF_OK;
  TDSSF_REQUIRE((((uintptr_t)x_pt | (uintptr_t)x_img | (uintptr_t)x_bf16 | (uintptr_t)h1_bf16 | (uintptr_t)h2_bf16 | (uintptr_t)dlogits |
                  (uintptr_t)dx_pt) & 31) == 0, "tdssf_mlp_bwd_tc: pointers must be 32-byte aligned");
  TDSSF_REQUIRE(workspace && ((uintptr_t)workspace & 255) == 0 && workspace_bytes >= tdssf_mlp_bwd_tc_workspace_bytes(N),
                "tdssf_mlp_bwd_tc: 256-byte aligned workspace of %zu bytes needed, got %zu",
                tdssf_mlp_bwd_tc_workspace_bytes(N), workspace_bytes);
  if (int rc = tdssf_check_device()) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const int n_cta = wgrad_ctas(N);
  Work w;
  carve(N, n_cta, reinterpret_cast<char*>(workspace), &w);
  const unsigned blocks = (unsigned)num_sms() * 8;

  pack_chain_weights_kernel<<<64, 256, 0, s>>>(w_host->W1, w_host->W2, w_host->W3, w.chain_w);
  TDSSF_LAUNCH_CHECK();
  const void* X = x_bf16;
  if (!X) {
    concat_bf16_kernel<<<blocks, 256, 0, s>>>(x_pt, x_img, reinterpret_cast<uint4*>(w.X), N);
    TDSSF_LAUNCH_CHECK();
    X = w.X;
  }
  CUtensorMap tm_h1_128, tm_dh1, tm_x, tm_h1, tm_dh2;
  if (!tmap::make_map(&tm_h1_128, h1_bf16, N, kN1, 128) || !tmap::make_map(&tm_dh1, w.dh1, N, kN1, wg::kKP) || !tmap::make_map(&tm_x, X, N, kK1, wg::kKP) ||
      !tmap::make_map(&tm_h1, h1_bf16, N, kN1, wg::kKP) || !tmap::make_map(&tm_dh2, w.dh2, N, kN2, wg::kKP)) {
    set_error("tdssf_mlp_bwd_tc: cuTensorMapEncodeTiled failed (driver entry point unavailable or bad tensor geometry)");
    return TDSSF_ECUDA;
  }

  // 1. data-gradient chain (writes dh2, dh1, dx_pt and per-CTA bias sums)
  int n_chain = 0;
  {
    static bool done[kMaxDevices] = {};
    if (int rc = set_smem(chain::dgrad_chain_kernel, chain::kSmemBytes + 1024, done)) return rc;
    chain::Params p = {};
    p.dl = dlogits; p.h2 = reinterpret_cast<const uint16_t*>(h2_bf16); p.weights = w.chain_w;
    p.dh2_out = w.dh2; p.dh1_out = w.dh1; p.dx_pt = dx_pt; p.bias_partial = w.bias_partials;
    p.N = N; p.n_tiles = (int)ceil_div(N, 128);
    const int grid = std::min(num_sms(), p.n_tiles);
    n_chain = grid;
#ifdef TDSSF_CH_TRACE
    static long long* trace = nullptr;
    if (!trace) { cudaMalloc(&trace, 17 * 16 * 8); }
    cudaMemsetAsync(trace, 0, 17 * 16 * 8, s);
    p.trace = trace;
#endif
    chain::dgrad_chain_kernel<<<grid, chain::kThreads, chain::kSmemBytes + 1024, s>>>(tm_h1_128, p);
    TDSSF_LAUNCH_CHECK();
#ifdef TDSSF_CH_TRACE
    {
      static int calls = 0;
      if (++calls == 6) {
        long long h[17 * 16];
        cudaStreamSynchronize(s);
        cudaMemcpy(h, trace, sizeof(h), cudaMemcpyDeviceToHost);
        const char* names[17] = {"epi acc1 full", "epi E1 done", "epi accx full", "epi Ex done", "prod start", "prod computed", "prod slab free",
                                 "prod arrived", "mma dh2+acc1free", "mma L1 issued", "mma dh1 slab0", "mma Lx issued", "help dh2 full",
                                 "help dh2 drained", "help dh1 slab0", "help dh1 drained", "tma issue"};
        long long t0 = h[4 * 16];
        fprintf(stderr, "[chain trace] CTA 0, cycles since the producer's first start; tiles 0..7\n");
        for (int e = 0; e < 17; ++e) {
          fprintf(stderr, "  %-18s", names[e]);
          for (int it = 0; it < 8; ++it) fprintf(stderr, " %8lld", h[e * 16 + it] ? h[e * 16 + it] - t0 : -1);
          fprintf(stderr, "\n");
        }
      }
    }
#endif
  }
  // 2. dW3, db3 on the CUDA cores
  const int n3 = (int)std::max<int64_t>(1, std::min<int64_t>(std::min(num_sms() * 2, kMaxW3Blocks), ceil_div(N, 256)));
  dw3_db3_kernel<<<n3, 256, 0, s>>>(dlogits, reinterpret_cast<const uint16_t*>(h2_bf16), N, w.partials3);
  TDSSF_LAUNCH_CHECK();
  // 3. dW1, dW2: per-CTA partials in TMEM, then one ordered reduction into the caller's buffers
  {
    static bool done[kMaxDevices] = {};
    if (int rc = set_smem(wg::wgrad_kernel, wg::kSmemBytes + 1024, done)) return rc;
    wg::Params p = {};
    p.partials = w.partials; p.N = N; p.n_stages = (int)ceil_div(N, wg::kKP);
    p.n_a = std::max(1, std::min(n_cta - 1, (n_cta * 2 + 2) / 5));   // bytes per stage 8 : 12 slabs -> CTAs 2 : 3
    wg::wgrad_kernel<<<n_cta, wg::kThreads, wg::kSmemBytes + 1024, s>>>(tm_dh1, tm_x, tm_h1, tm_dh2, p);
    TDSSF_LAUNCH_CHECK();
    wg::wgrad_reduce_kernel<<<(kN1 * kK1 + kN1 * kN2 + 255) / 256 + (kW3Out + 31) / 32 + 10, 256, 0, s>>>(
        w.partials, p.n_a, n_cta, g_host->dW1, g_host->dW2, w.partials3, n3, g_host->dW3, g_host->db3, w.bias_partials, n_chain,
        g_host->db1, g_host->db2);
    TDSSF_LAUNCH_CHECK();
  }
  return TDSSF_OK;
}
