// a1..a6 fused inference on the 5th-generation tensor cores (tcgen05, bf16 operands, fp32 accumulate in
// TMEM):   project -> mask -> patch gather -> masked camera mean -> concat -> 448->256->64->16 MLP -> logits
// Replaces FeatureFusionModel.forward up to voxel_scores (utils/fusion_model.py:152-204).
//
// One persistent 2-CTA cluster per SM pair (cta_group::2).  Each CTA owns a 128-point tile; the pair
// shares every B operand: W1 (256x448 bf16 = 224 KB) does not fit one SM's shared memory, split in two
// N-halves of 112 KB it stays resident for the whole kernel, and so do W2 and W3.
//
//   warp 0        TMEM allocation, weight load (cp.async.bulk), MMA issue (leader CTA only)
//   warps 1-4     epilogue: TMEM -> registers -> bias/ReLU -> bf16 -> TMEM (A operand of the next layer);
//                 last layer -> logits to HBM
//   warps 5-20    producers: projection of the tile's points (records in smem), then per 64-channel K slab:
//                 gather + camera mean from the L2-resident token maps (or dense loads), bf16, written in
//                 the 128B-swizzled K-major layout the UMMA descriptor expects
//
// TMEM (512 columns) holds two 256-column tile buffers; inside a buffer the layer-1 accumulator is
// overwritten in place by its own bf16 activation (H1), and the layer-2/3 accumulators and H2 reuse
// columns of it that have already been consumed:
//   [0,256) acc1   ->   [0,128) H1 | [128,192) acc2 | [192,224) H2 | [224,256) acc3
// MMA issue order  L1(t+1), L2(t), L3(t)  overlaps the layer-1 epilogue of tile t with the layer-1
// MMAs of tile t+1.
#include <algorithm>
#include "project.cuh"
#include "tc_ptx.cuh"

namespace tdssf {
namespace tc {

constexpr int kTileM = 128;
constexpr int kCp = 64, kCi = 384, kK1 = kCp + kCi;  // 448
constexpr int kN1 = 256, kN2 = 64, kN3 = 32 /* 16 classes padded to 32 */, kClasses = 16;
constexpr int kSlabK = 64;                           // K elements per 128-byte swizzled row
constexpr int kSlabs1 = kK1 / kSlabK;                // 7
constexpr int kSlabs2 = kN1 / kSlabK;                // 4
constexpr int kStages = 5;
constexpr int kProducerWarps = 16, kEpiWarps = 4;
constexpr int kThreads = 32 * (1 + kEpiWarps + kProducerWarps);  // 672
constexpr int kRowsPerWarp = kTileM / kProducerWarps;            // 8

// per-CTA shared-memory image (bytes)
constexpr int kW1Slab = (kN1 / 2) * 128;   // 128 rows x 128 B = 16 KB
constexpr int kW2Slab = (kN2 / 2) * 128;   // 32 rows  -> 4 KB
constexpr int kW3Slab = (kN3 / 2) * 128;   // 16 rows  -> 2 KB
constexpr int kOffW1 = 0;
constexpr int kOffW2 = kOffW1 + kSlabs1 * kW1Slab;          // 114688
constexpr int kOffW3 = kOffW2 + kSlabs2 * kW2Slab;          // 131072
constexpr int kWeightBytes = kOffW3 + kW3Slab;              // 133120 per CTA rank
constexpr int kBiasFloats = kN1 + kN2 + kClasses;           // 336
constexpr int kOffRing = kWeightBytes;                      // 1024-aligned
constexpr int kASlab = kTileM * 128;                        // 16 KB
constexpr int kOffBias = kOffRing + kStages * kASlab;       // 215040
constexpr int kOffRec = kOffBias + ((kBiasFloats * 4 + 15) & ~15);
constexpr int kOffBar = kOffRec + 2 * kTileM * 32;     // [warp][2 tiles][8 rows] x 8 slot words
constexpr int kNumBars = 2 * kStages + 2 + 6 * 2;           // a_full, a_empty, w_full, peer_w, 6 per-buffer kinds x 2
constexpr int kOffTmemPtr = kOffBar + kNumBars * 8;
constexpr int kSmemBytes = kOffTmemPtr + 16;
static_assert(kSmemBytes + 1024 <= 232448, "shared memory budget exceeded");
static_assert(kOffRing % 1024 == 0 && kOffW2 % 1024 == 0 && kOffW3 % 1024 == 0, "slab alignment");

constexpr size_t kPackBytes = 2 * (size_t)kWeightBytes + kBiasFloats * 4 + 64;

// barrier indices
enum {
  kBarAFull = 0,                       // [kStages]   leader: 2*kProducerWarps arrivals per slab
  kBarAEmpty = kStages,                // [kStages]   each CTA: 1 arrival (commit multicast)
  kBarWFull = 2 * kStages,             // weights landed (tx bytes)
  kBarPeerW = kBarWFull + 1,           // leader: the peer CTA's weights landed (1 remote arrival)
  kBarAcc1Full = kBarPeerW + 1,        // [2] each CTA, commit
  kBarAcc2Full = kBarAcc1Full + 2,     // [2]
  kBarAcc3Full = kBarAcc2Full + 2,     // [2]
  kBarH1Full = kBarAcc3Full + 2,       // [2] leader: 2*kEpiWarps arrivals
  kBarH2Full = kBarH1Full + 2,         // [2]
  kBarAccFree = kBarH2Full + 2,        // [2]
};

struct Params {
  // source of the 384 image channels: gather (tokens + projection or records) or a dense matrix
  const float* xyz; int64_t sweep_stride, point_stride;
  const float* K; const float* Rt; const float* img_hw;
  int V, NC, resize_h, resize_w, patch, M;
  const void* tokens;
  const tdssf_rec* rec_in;     // optional precomputed records (skips the projection)
  tdssf_rec* rec_out;          // optional
  const float* x_img;          // dense mode
  const float* x_pt;           // [N,64]
  const uint8_t* packed;       // weights image (tdssf_mlp_pack_weights)
  float* logits;               // [N,16]
  int64_t N;
  int n_pairs;                 // number of 256-row tile pairs
  unsigned int* debug;         // optional: [0] = timeout code
};

// ---- optional cycle accounting (experiment builds only: -DTDSSF_PROFILE) -----------------------------------
#ifdef TDSSF_PROFILE
#define PROF_T0(var) const long long var = clock64()
#define PROF_ADD(dbg, code, var)                                                                        \
  do {                                                                                                  \
    if ((threadIdx.x & 31) == 0 && (dbg))                                                               \
      atomicAdd(reinterpret_cast<unsigned long long*>(dbg) + 2 + (code), (unsigned long long)(clock64() - var)); \
  } while (0)
#else
#define PROF_T0(var)
#define PROF_ADD(dbg, code, var)
#endif

__device__ __forceinline__ void wait_bar(uint32_t bar, uint32_t parity, unsigned int* debug, int code) {
  if (ptx::mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
#ifdef TDSSF_PROFILE
  struct Acc { unsigned int* d; int c; long long t; __device__ ~Acc() { PROF_ADD(d, c, t); } } acc_{debug, code, t0};
#endif
  while (!ptx::mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {  // ~2 s: a protocol bug, never a legitimate wait
      if (debug) atomicExch(debug, (unsigned)code | 0x80000000u);
      __trap();
    }
  }
}


__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ---- producer loads ---------------------------------------------------------------------------------------
// A quarter-warp (8 lanes) owns one row of a 64-channel K slab.  fp32 sources (256 B per row piece) are read as two
// fully coalesced 128-byte requests: lane ql takes channels [4ql,4ql+4) and [32+4ql,32+4ql+4).  16-bit sources
// (128 B) are one request: lane ql takes channels [8ql,8ql+8).  Loads return raw bits; conversion is deferred to the
// consumer so that the request stays in flight.
#ifndef TDSSF_LDMODE
#define TDSSF_LDMODE 0  // 0: ld.global.nc   1: ld.global.cg (L2 only)   2: ld.global.nc.L1::no_allocate
#endif
__device__ __forceinline__ uint4 ldg16(const void* p) {
#if TDSSF_LDMODE == 1
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
#elif TDSSF_LDMODE == 2
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
#else
  return __ldg(reinterpret_cast<const uint4*>(p));
#endif
}
struct Raw8 { uint4 lo, hi; };
__device__ __forceinline__ Raw8 raw8_zero() { Raw8 r; r.lo = make_uint4(0, 0, 0, 0); r.hi = r.lo; return r; }
// piece = start of the row's 64-channel piece
__device__ __forceinline__ Raw8 raw8_load_f32(const float* piece, int ql) {
  Raw8 r;
  r.lo = ldg16(piece + ql * 4);
  r.hi = ldg16(piece + 32 + ql * 4);
  return r;
}
__device__ __forceinline__ void raw8_f32_to_f8(const Raw8& q, float (&f)[8]) {
  f[0] = __uint_as_float(q.lo.x); f[1] = __uint_as_float(q.lo.y); f[2] = __uint_as_float(q.lo.z); f[3] = __uint_as_float(q.lo.w);
  f[4] = __uint_as_float(q.hi.x); f[5] = __uint_as_float(q.hi.y); f[6] = __uint_as_float(q.hi.z); f[7] = __uint_as_float(q.hi.w);
}
template <typename T> struct TokRaw;
template <> struct TokRaw<float> {
  static constexpr bool kSplit = true;
  __device__ static Raw8 load(const float* piece, int ql) { return raw8_load_f32(piece, ql); }
  __device__ static void to_f8(const Raw8& q, float (&f)[8]) { raw8_f32_to_f8(q, f); }
};
template <> struct TokRaw<__half> {
  static constexpr bool kSplit = false;
  __device__ static Raw8 load(const __half* piece, int ql) {
    Raw8 r; r.lo = ldg16(piece + ql * 8); r.hi = make_uint4(0, 0, 0, 0); return r;
  }
  __device__ static void to_f8(const Raw8& q, float (&f)[8]) {
    uint32_t w[4] = {q.lo.x, q.lo.y, q.lo.z, q.lo.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 a = __half22float2(*reinterpret_cast<__half2*>(&w[i])); f[2 * i] = a.x; f[2 * i + 1] = a.y; }
  }
};
template <> struct TokRaw<__nv_bfloat16> {
  static constexpr bool kSplit = false;
  __device__ static Raw8 load(const __nv_bfloat16* piece, int ql) {
    Raw8 r; r.lo = ldg16(piece + ql * 8); r.hi = make_uint4(0, 0, 0, 0); return r;
  }
  __device__ static void to_f8(const Raw8& q, float (&f)[8]) {
    uint32_t w[4] = {q.lo.x, q.lo.y, q.lo.z, q.lo.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 a = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w[i])); f[2 * i] = a.x; f[2 * i + 1] = a.y; }
  }
};
// bf16 A-operand row r of a slab, 128B-swizzled: 16-byte chunk c lives at ((c ^ (r & 7)) << 4)
template <bool kSplit>
__device__ __forceinline__ void store_row8(uint8_t* slab, int r, int ql, const float (&f)[8]) {
  const uint32_t row_off = (uint32_t)r * 128u, sw = (uint32_t)r & 7u;
  if (kSplit) {  // channels [4ql,4ql+4) -> chunk ql>>1, half ql&1;  channels [32+4ql, ...) -> chunk 4 + (ql>>1)
    const uint32_t c0 = (uint32_t)ql >> 1, hb = ((uint32_t)ql & 1u) * 8u;
    *reinterpret_cast<uint2*>(slab + row_off + ((c0 ^ sw) << 4) + hb) = make_uint2(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]));
    *reinterpret_cast<uint2*>(slab + row_off + (((c0 + 4u) ^ sw) << 4) + hb) = make_uint2(pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
  } else {
    *reinterpret_cast<uint4*>(slab + row_off + ((((uint32_t)ql) ^ sw) << 4)) =
        make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
  }
}

// cameras beyond the preloaded ones (rare): sequential loads in camera order, straight from the slot words
template <typename TokT>
__device__ __noinline__ void gather_rest(const uint32_t* sl, int covered, const TokT* tokens, int NC, int M, int ch0,
                                         int ql, float (&acc)[8]) {
  const uint32_t tok_row_base = sl[6] * (uint32_t)(NC * M);
  int k = 0;
  for (int c = 0; c < TDSSF_MAX_CAMS; ++c) {
    const uint32_t v = sl[c];
    if (v >> 31) {
      if (k >= covered) {
        float f[8];
        TokRaw<TokT>::to_f8(TokRaw<TokT>::load(tokens + (int64_t)(tok_row_base + c * M + (v & 0xffffu)) * kCi + ch0, ql), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], f[i]);
      }
      k++;
    }
  }
}

// kGather: image channels come from the token maps through the records; else from p.x_img
template <typename TokT, bool kGather>
__global__ void __launch_bounds__(kThreads, 1)
fused_mlp_kernel(const Params p) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the 128B-swizzled operand slabs
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int n_iters = (p.n_pairs - (int)ptx::cluster_id_x() + (int)ptx::num_clusters_x() - 1) / (int)ptx::num_clusters_x();
  auto bar = [&](int i) { return sbase + kOffBar + 8 * i; };
  float* s_bias = reinterpret_cast<float*>(smem + kOffBias);
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + kOffTmemPtr);

  // ---- one-time setup ------------------------------------------------------------------------------
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(bar(kBarAFull + s), 2 * kProducerWarps);
      ptx::mbar_init(bar(kBarAEmpty + s), 1);
    }
    ptx::mbar_init(bar(kBarWFull), 1);
    ptx::mbar_init(bar(kBarPeerW), 1);
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(bar(kBarAcc1Full + b), 1);
      ptx::mbar_init(bar(kBarAcc2Full + b), 1);
      ptx::mbar_init(bar(kBarAcc3Full + b), 1);
      ptx::mbar_init(bar(kBarH1Full + b), 2 * kEpiWarps);
      ptx::mbar_init(bar(kBarH2Full + b), 2 * kEpiWarps);
      ptx::mbar_init(bar(kBarAccFree + b), 2 * kEpiWarps);
    }
    ptx::fence_mbar_init();
  }
  {  // biases (shared by both ranks) live after the two weight images
    const float* gb = reinterpret_cast<const float*>(p.packed + 2 * (size_t)kWeightBytes);
    for (int i = threadIdx.x; i < kBiasFloats; i += kThreads) s_bias[i] = __ldg(gb + i);
  }
  if (warp == 0) {
    ptx::tmem_alloc_2cta(sbase + kOffTmemPtr, 512);
    ptx::tmem_relinquish_2cta();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();  // barriers initialised and TMEM allocated in both CTAs before any remote traffic
  ptx::tc_fence_after();
  const uint32_t tmem = *s_tmem;

  if (warp == 0) {
    // =============================== weights + MMA issue ==========================================
    if (lane == 0) {
      const uint8_t* src = p.packed + (size_t)rank * kWeightBytes;
      ptx::mbar_expect_tx(bar(kBarWFull), kWeightBytes);
      for (int off = 0; off < kWeightBytes; off += 16384) {
        const int n = min(16384, kWeightBytes - off);
        ptx::bulk_g2s(sbase + kOffW1 + off, src + off, n, bar(kBarWFull));
      }
    }
    if (leader && lane == 0 && n_iters > 0) {
      // B operands are read from both CTAs' shared memory: wait for our image and for the peer's
      wait_bar(bar(kBarWFull), 0, p.debug, 1);
      wait_bar(bar(kBarPeerW), 0, p.debug, 11);
      constexpr uint32_t kI1 = ptx::idesc_bf16(256, kN1), kI2 = ptx::idesc_bf16(256, kN2), kI3 = ptx::idesc_bf16(256, kN3);
      uint32_t g = 0;  // global slab counter
      auto issue_l1 = [&](int it) {
        const uint32_t buf = it & 1;
        if (it >= 2) wait_bar(bar(kBarAccFree + buf), ((it >> 1) - 1) & 1, p.debug, 2);
        ptx::tc_fence_after();
        for (int s = 0; s < kSlabs1; ++s, ++g) {
          const uint32_t stage = g % kStages, par = (g / kStages) & 1;
          wait_bar(bar(kBarAFull + stage), par, p.debug, 3);
          ptx::tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const uint64_t ad = ptx::smem_desc_sw128(sbase + kOffRing + stage * kASlab + kk * 32);
            const uint64_t bd = ptx::smem_desc_sw128(sbase + kOffW1 + s * kW1Slab + kk * 32);
            ptx::umma_ss_2cta(tmem + buf * 256, ad, bd, kI1, (s | kk) != 0);
          }
          ptx::umma_commit_2cta(bar(kBarAEmpty + stage), 3);
        }
        ptx::umma_commit_2cta(bar(kBarAcc1Full + buf), 3);
      };
      PROF_T0(t25);
      issue_l1(0);
      for (int it = 0; it < n_iters; ++it) {
        const uint32_t buf = it & 1, par = (it >> 1) & 1;
        if (it + 1 < n_iters) issue_l1(it + 1);
        // layer 2: A = H1 (TMEM cols [0,128) of the buffer), B = W2
        wait_bar(bar(kBarH1Full + buf), par, p.debug, 4);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < kN1 / 16; ++k) {
          const uint64_t bd = ptx::smem_desc_sw128(sbase + kOffW2 + (k >> 2) * kW2Slab + (k & 3) * 32);
          ptx::umma_ts_2cta(tmem + buf * 256 + 128, tmem + buf * 256 + k * 8, bd, kI2, k != 0);
        }
        ptx::umma_commit_2cta(bar(kBarAcc2Full + buf), 3);
        // layer 3: A = H2 (cols [192,224)), B = W3
        wait_bar(bar(kBarH2Full + buf), par, p.debug, 5);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < kN2 / 16; ++k) {
          const uint64_t bd = ptx::smem_desc_sw128(sbase + kOffW3 + k * 32);
          ptx::umma_ts_2cta(tmem + buf * 256 + 224, tmem + buf * 256 + 192 + k * 8, bd, kI3, k != 0);
        }
        ptx::umma_commit_2cta(bar(kBarAcc3Full + buf), 3);
      }
      PROF_ADD(p.debug, 25, t25);
    }
    if (!leader && lane == 0) {
      wait_bar(bar(kBarWFull), 0, p.debug, 6);
      ptx::mbar_arrive_cluster(ptx::mapa(bar(kBarPeerW), 0));
    }
    __syncwarp();
  } else if (warp <= kEpiWarps) {
    // =============================== epilogue ======================================================
    const uint32_t lane_base = (uint32_t)(warp & 3) * 32;
    const uint32_t tlane = tmem + (lane_base << 16);
    const int row_in_tile = (int)lane_base + lane;
    const uint32_t leader_h1 = ptx::mapa(bar(kBarH1Full), 0), leader_h2 = ptx::mapa(bar(kBarH2Full), 0),
                   leader_free = ptx::mapa(bar(kBarAccFree), 0);
    PROF_T0(t26);
    for (int it = 0; it < n_iters; ++it) {
      const uint32_t buf = it & 1, par = (it >> 1) & 1;
      const int64_t pair = (int64_t)ptx::cluster_id_x() + (int64_t)it * ptx::num_clusters_x();
      const int64_t row = (pair * 2 + rank) * kTileM + row_in_tile;
      const uint32_t tb = tlane + buf * 256;
      // ---- layer 1: acc1 -> relu(+b1) -> bf16 H1, in place --------------------------------------
      wait_bar(bar(kBarAcc1Full + buf), par, p.debug, 7);
      ptx::tc_fence_after();
      PROF_T0(t27);
#pragma unroll 1
      for (int q = 0; q < kN1 / 32; ++q) {
        uint32_t v[32], h[16];
        ptx::tmem_ld32(tb + q * 32, v);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float a = fmaxf(__uint_as_float(v[2 * j]) + s_bias[q * 32 + 2 * j], 0.f);
          const float b = fmaxf(__uint_as_float(v[2 * j + 1]) + s_bias[q * 32 + 2 * j + 1], 0.f);
          h[j] = pack_bf16(a, b);
        }
        ptx::tmem_st16(tb + q * 16, h);
      }
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(leader_h1 + 8 * buf);
      __syncwarp();
      PROF_ADD(p.debug, 27, t27);
      // ---- layer 2: acc2 -> relu(+b2) -> bf16 H2 --------------------------------------------------
      wait_bar(bar(kBarAcc2Full + buf), par, p.debug, 8);
      ptx::tc_fence_after();
#pragma unroll
      for (int q = 0; q < kN2 / 32; ++q) {
        uint32_t v[32], h[16];
        ptx::tmem_ld32(tb + 128 + q * 32, v);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float a = fmaxf(__uint_as_float(v[2 * j]) + s_bias[kN1 + q * 32 + 2 * j], 0.f);
          const float b = fmaxf(__uint_as_float(v[2 * j + 1]) + s_bias[kN1 + q * 32 + 2 * j + 1], 0.f);
          h[j] = pack_bf16(a, b);
        }
        ptx::tmem_st16(tb + 192 + q * 16, h);
      }
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(leader_h2 + 8 * buf);
      __syncwarp();
      // ---- layer 3: acc3 + b3 -> logits ------------------------------------------------------------
      wait_bar(bar(kBarAcc3Full + buf), par, p.debug, 9);
      ptx::tc_fence_after();
      {
        uint32_t v[16];
        ptx::tmem_ld16(tb + 224, v);
        ptx::tmem_wait_ld();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_cluster(leader_free + 8 * buf);
        __syncwarp();
        if (row < p.N) {
          float4* dst = reinterpret_cast<float4*>(p.logits + row * kClasses);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            dst[j] = make_float4(__uint_as_float(v[4 * j]) + s_bias[kN1 + kN2 + 4 * j],
                                 __uint_as_float(v[4 * j + 1]) + s_bias[kN1 + kN2 + 4 * j + 1],
                                 __uint_as_float(v[4 * j + 2]) + s_bias[kN1 + kN2 + 4 * j + 2],
                                 __uint_as_float(v[4 * j + 3]) + s_bias[kN1 + kN2 + 4 * j + 3]);
        }
      }
    }
    PROF_ADD(p.debug, 26, t26);
  } else {
    // =============================== producers =====================================================
    // Per tile: (1) every (row, camera) pair is projected by its own thread into a slot word
    // "valid<<31 | patch" (done one tile ahead, while gather loads are in flight); (2) each warp compacts
    // the slots of its 8 rows into "first / second valid token row"; (3) per 64-channel K slab all loads of
    // the warp's rows are issued before any is consumed, and the next slab's loads are issued before the
    // current slab is converted (two register sets) so that ~2 slabs of L2 requests stay in flight per SM.
    const int pw = warp - 1 - kEpiWarps;                  // 0..15
    const int ptid = threadIdx.x - 32 * (1 + kEpiWarps);  // 0..511
    const int quarter = lane >> 3, ql = lane & 7;
    const uint32_t leader_full0 = ptx::mapa(bar(kBarAFull), 0);
    uint32_t* s_slot = reinterpret_cast<uint32_t*>(smem + kOffRec);  // [16 warps][2 tiles][8 rows][8]
    auto tile_row0 = [&](int it) {
      return (((int64_t)ptx::cluster_id_x() + (int64_t)it * ptx::num_clusters_x()) * 2 + rank) * kTileM;
    };

    // warp-private slot words: [warp][2 tiles][8 rows][8] — each warp projects, compacts and gathers its own 8 rows,
    // so the producers never wait for one another
    auto slots_of = [&](int it) { return s_slot + ((pw * 2 + (it & 1)) * kRowsPerWarp) * 8; };
    auto fill_slots = [&](int it) {
      uint32_t* slots = slots_of(it);
      const int64_t row0 = tile_row0(it) + pw * kRowsPerWarp;
      if (p.rec_in) {
        if (lane < kRowsPerWarp) {
          const int64_t row = row0 + lane;
          uint4 r = make_uint4(0, 0, 0, 0);
          uint32_t b = 0;
          if (row < p.N) {
            r = __ldg(reinterpret_cast<const uint4*>(p.rec_in) + row);
            b = (uint32_t)(row / p.V);
          }
          const uint32_t pk[3] = {r.x, r.y, r.z};
#pragma unroll
          for (int c = 0; c < TDSSF_MAX_CAMS; ++c)
            slots[lane * 8 + c] = ((pk[c >> 1] >> ((c & 1) * 16)) & 0xffffu) | (((r.w >> c) & 1u) << 31);
          slots[lane * 8 + 6] = b;
        }
        return;
      }
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const int j = lane + 32 * jj;  // 48 (row, camera) pairs over 32 lanes
        if (j < kRowsPerWarp * TDSSF_MAX_CAMS) {
          const int cam = j >> 3, r = j & (kRowsPerWarp - 1);
          const int64_t row = row0 + r;
          uint32_t val = 0, b = 0;
          if (row < p.N) {
            b = (uint32_t)(row / p.V);
            if (cam < p.NC) {
              const int v = (int)(row - (int64_t)b * p.V);
              const float* q = p.xyz + b * p.sweep_stride + v * p.point_stride;
              const float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
              const SweepGeom geo = make_geom(__ldg(p.img_hw + 2 * b), __ldg(p.img_hw + 2 * b + 1), p.resize_h, p.resize_w, p.patch);
              CamCalib cal;
              const float* rt = p.Rt + ((int64_t)b * p.NC + cam) * 16;
              const float* kk = p.K + ((int64_t)b * p.NC + cam) * 9;
#pragma unroll
              for (int i = 0; i < 12; ++i) (&cal.r[0][0])[i] = __ldg(rt + i);
#pragma unroll
              for (int i = 0; i < 9; ++i) (&cal.k[0][0])[i] = __ldg(kk + i);
              const Proj pr = project_one(cal, x, y, z, geo.img_h, geo.img_w);
              val = (uint32_t)patch_index(geo, pr.u, pr.v) | (pr.valid ? 0x80000000u : 0u);
            }
          }
          slots[r * 8 + cam] = val;
          if (cam == 0) slots[r * 8 + 6] = b;
        }
      }
    };

    // L2 prefetch (TMA bulk prefetch, no registers, no completion tracking).  The gather visits every token row six
    // times, slab by slab, so demand misses would hit DRAM with scattered 256-byte reads at full DRAM latency; instead
    // the whole token map of the sweeps about to be processed is streamed into L2 sequentially, each CTA taking every
    // gridDim-th 16 KB chunk, and the next tiles' point features / points are prefetched as they are contiguous.
    const int64_t tok_sweep_bytes = (int64_t)p.NC * p.M * kCi * (int64_t)sizeof(TokT);
    const int pf_ahead = tok_sweep_bytes <= (32 << 20) ? 2 : 1;
    int pf_next = 0;  // first sweep whose token map has not been requested yet (meaningful in thread ptid == 0)
    auto prefetch_for_tile = [&](int it) {
      if (ptid != 0) return;
      const int64_t row0 = tile_row0(it);
      if (kGather) {
        const int n_sweeps = (int)((p.N + p.V - 1) / p.V);
        const int want = min(n_sweeps, (int)(min((long long)(row0 + kTileM - 1), (long long)(p.N - 1)) / p.V) + 1 + pf_ahead);
        for (; pf_next < want; ++pf_next) {
          const uint8_t* base = reinterpret_cast<const uint8_t*>(p.tokens) + (int64_t)pf_next * tok_sweep_bytes;
          for (int64_t off = (int64_t)blockIdx.x * 16384; off < tok_sweep_bytes; off += (int64_t)gridDim.x * 16384)
            ptx::prefetch_l2(base + off, (uint32_t)min((long long)16384, (long long)(tok_sweep_bytes - off)));
        }
      }
      const int64_t r2 = tile_row0(it + 2);  // point features (and points) two tiles ahead: contiguous blocks
      if (it + 2 < n_iters && r2 < p.N) {
        const int64_t rows = min((long long)kTileM, (long long)(p.N - r2));
        ptx::prefetch_l2(p.x_pt + r2 * kCp, (uint32_t)(rows * kCp * 4));
        if (!kGather) {
          ptx::prefetch_l2(p.x_img + r2 * kCi, (uint32_t)(rows * kCi * 4));
        } else if (!p.rec_in && r2 / p.V == (r2 + rows - 1) / p.V && (p.point_stride & 3) == 0) {
          const int64_t b = r2 / p.V;
          const float* q = p.xyz + b * p.sweep_stride + (r2 - b * p.V) * p.point_stride;
          if ((reinterpret_cast<uintptr_t>(q) & 15) == 0) ptx::prefetch_l2(q, (uint32_t)(rows * p.point_stride * 4));
        }
      }
    };

    // quarter-warp q owns row pw*8 + step*4 + q of the tile; lane ql of the quarter owns 8 channels of the slab
    constexpr int kSteps = kRowsPerWarp / 4;  // 2
    struct TileRec {
      uint32_t base0[kSteps];  // token row (sweep, camera, patch) of the first valid camera
      uint32_t cn[kSteps];     // number of valid cameras
      uint32_t ovb;            // token row of the second camera of step `ovs`
      int ovs;                 // the step (if any) whose row has a second camera preloaded, else -1
    };

    // compaction of the slots of tile `it`: lane j < 8 owns row pw*8 + j, results are shuffled to the quarters
    auto compact = [&](int it, TileRec& R) {
      const uint32_t* slots = slots_of(it);
      const int64_t row0 = tile_row0(it);
      uint32_t c_i0 = 0, c_i1 = 0, c_cnt = 0;
      if (lane < kRowsPerWarp) {
        const int r = pw * kRowsPerWarp + lane;
        const uint4 a = *reinterpret_cast<const uint4*>(slots + lane * 8);
        const uint4 bq = *reinterpret_cast<const uint4*>(slots + lane * 8 + 4);
        const uint32_t sl[6] = {a.x, a.y, a.z, a.w, bq.x, bq.y};
        const uint32_t tok_row_base = bq.z * (uint32_t)(p.NC * p.M);
        uint32_t mask = 0;
#pragma unroll
        for (int c = 0; c < TDSSF_MAX_CAMS; ++c) {
          if (sl[c] >> 31) {
            const uint32_t idx = tok_row_base + (uint32_t)c * p.M + (sl[c] & 0xffffu);
            if (c_cnt == 0) c_i0 = idx;
            else if (c_cnt == 1) c_i1 = idx;
            c_cnt++;
            mask |= 1u << c;
          }
        }
        if (p.rec_out && !p.rec_in && row0 + r < p.N)
          reinterpret_cast<uint4*>(p.rec_out)[row0 + r] =
              make_uint4((sl[0] & 0xffffu) | (sl[1] << 16), (sl[2] & 0xffffu) | (sl[3] << 16),
                         (sl[4] & 0xffffu) | (sl[5] << 16), mask | (c_cnt << 8));
      }
      R.ovs = -1;
      R.ovb = 0;
#pragma unroll
      for (int step = 0; step < kSteps; ++step) {
        const int src = step * 4 + quarter;
        R.base0[step] = __shfl_sync(0xffffffffu, c_i0, src);
        const uint32_t i1 = __shfl_sync(0xffffffffu, c_i1, src);
        R.cn[step] = __shfl_sync(0xffffffffu, c_cnt, src);
        if (R.cn[step] >= 2 && R.ovs < 0) { R.ovs = step; R.ovb = i1; }
      }
    };

    auto tok_piece = [&](uint32_t tok_row, int s) {
      return reinterpret_cast<const TokT*>(p.tokens) + (int64_t)tok_row * kCi + (s - 1) * kSlabK;
    };
    // loads of slab s of the tile whose first row is row0 (records R); nothing is consumed here
    auto issue = [&](int64_t row0, const TileRec& R, int s, Raw8 (&t)[kSteps], Raw8& u) {
#pragma unroll
      for (int step = 0; step < kSteps; ++step) {
        const int64_t row = row0 + pw * kRowsPerWarp + step * 4 + quarter;
        t[step] = raw8_zero();
        if (s == 0) {
          if (row < p.N) t[step] = raw8_load_f32(p.x_pt + row * kCp, ql);
        } else if (!kGather) {
          if (row < p.N) t[step] = raw8_load_f32(p.x_img + row * kCi + (s - 1) * kSlabK, ql);
        } else if (R.cn[step] >= 1) {
          t[step] = TokRaw<TokT>::load(tok_piece(R.base0[step], s), ql);
        }
      }
      if (kGather && s > 0 && R.ovs >= 0) u = TokRaw<TokT>::load(tok_piece(R.ovb, s), ql);
    };
    uint32_t g = 0;
    auto consume = [&](int it, const TileRec& R, int s, const Raw8 (&t)[kSteps], const Raw8& u) {
      const uint32_t stage = g % kStages, par = (g / kStages) & 1;
      wait_bar(bar(kBarAEmpty + stage), par ^ 1, p.debug, 10);
      uint8_t* slab = smem + kOffRing + stage * kASlab;
#pragma unroll
      for (int step = 0; step < kSteps; ++step) {
        const int r = pw * kRowsPerWarp + step * 4 + quarter;
        float acc[8];
        if (s == 0 || !kGather) {
          raw8_f32_to_f8(t[step], acc);
          store_row8<true>(slab, r, ql, acc);
        } else {
          TokRaw<TokT>::to_f8(t[step], acc);
          if (R.cn[step] >= 2) {  // camera-order sum, then one divide (utils/fusion_model.py:198-201)
            int covered = 1;
            if (step == R.ovs) {
              float f[8];
              TokRaw<TokT>::to_f8(u, f);
#pragma unroll
              for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], f[i]);
              covered = 2;
            }
            if ((int)R.cn[step] > covered)  // rare: a third camera, or two two-camera rows in one quarter-warp
              gather_rest<TokT>(slots_of(it) + (step * 4 + quarter) * 8, covered, reinterpret_cast<const TokT*>(p.tokens), p.NC, p.M,
                                (s - 1) * kSlabK, ql, acc);
            const float den = (float)R.cn[step];
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] = __fdiv_rn(acc[i], den);
          }
          store_row8<TokRaw<TokT>::kSplit>(slab, r, ql, acc);
        }
      }
#ifndef TDSSF_EXP_NOFENCE
      ptx::fence_proxy_async_smem();
#endif
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(leader_full0 + 8 * stage);
      __syncwarp();
      ++g;
    };

    // One tile.  On entry slab 0 (point features) of tile `it` is already in flight in A; on exit slab 0 of tile
    // it+1 is in flight in B, so the request stream never drains at a tile boundary.  The next tile's points are
    // projected at the start of this tile and compacted near its end (one named barrier in between).
    TileRec cur, nxt;
    auto tile = [&](int it, Raw8 (&A)[kSteps], Raw8& uA, Raw8 (&B)[kSteps], Raw8& uB) {
      const int64_t row0 = tile_row0(it);
      const bool more = it + 1 < n_iters;
#pragma unroll
      for (int s = 0; s < kSlabs1; ++s) {
        Raw8 (&tc)[kSteps] = (s & 1) ? B : A;   // holds slab s
        Raw8& uc = (s & 1) ? uB : uA;
        Raw8 (&tn)[kSteps] = (s & 1) ? A : B;   // receives slab s+1
        Raw8& un = (s & 1) ? uA : uB;
        { PROF_T0(t21);
        if (s + 1 < kSlabs1) issue(row0, cur, s + 1, tn, un);
        else if (more) issue(tile_row0(it + 1), cur, 0, tn, un);
        PROF_ADD(p.debug, 21, t21); }
        if (s == 0) prefetch_for_tile(it);
        if (kGather && more && s == 0) {  // next tile's projection + compaction hide under the loads in flight
          PROF_T0(t22);
          fill_slots(it + 1);
          __syncwarp();
          compact(it + 1, nxt);
          PROF_ADD(p.debug, 22, t22);
        }
        { PROF_T0(t20); consume(it, cur, s, tc, uc); PROF_ADD(p.debug, 20, t20); }
      }
      if (kGather && more) cur = nxt;
    };

    Raw8 ta[kSteps], tb[kSteps];
    Raw8 ua = raw8_zero(), ub = raw8_zero();
#pragma unroll
    for (int step = 0; step < kSteps; ++step) { cur.base0[step] = 0; cur.cn[step] = 0; }
    cur.ovb = 0; cur.ovs = -1;
    nxt = cur;
    if (n_iters > 0) {
      if (kGather) {
        fill_slots(0);
        __syncwarp();
        compact(0, cur);
      }
      issue(tile_row0(0), cur, 0, ta, ua);
      PROF_T0(t24);
      for (int it = 0; it < n_iters; it += 2) {
        tile(it, ta, ua, tb, ub);
        if (it + 1 < n_iters) tile(it + 1, tb, ub, ta, ua);
      }
      PROF_ADD(p.debug, 24, t24);
    }
  }

  // ---- teardown ----------------------------------------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 0) ptx::tmem_dealloc_2cta(tmem, 512);
}

// ---- weight packing: fp32 nn.Linear weights -> bf16 shared-memory images of both CTA ranks ------------------
__global__ void pack_weights_kernel(const float* __restrict__ W1, const float* __restrict__ b1,
                                    const float* __restrict__ W2, const float* __restrict__ b2,
                                    const float* __restrict__ W3, const float* __restrict__ b3,
                                    uint8_t* __restrict__ out) {
  const int total = 2 * (kWeightBytes / 2);
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int rank = e / (kWeightBytes / 2);
    const int byte = (e % (kWeightBytes / 2)) * 2;
    // invert the layout: which (layer, local row n, k) lives at this byte of the image?
    const float* W; int ld, rows_cta, off, n_valid;
    if (byte < kOffW2) { W = W1; ld = kK1; rows_cta = kN1 / 2; off = byte - kOffW1; n_valid = kN1; }
    else if (byte < kOffW3) { W = W2; ld = kN1; rows_cta = kN2 / 2; off = byte - kOffW2; n_valid = kN2; }
    else { W = W3; ld = kN2; rows_cta = kN3 / 2; off = byte - kOffW3; n_valid = kClasses; }
    const int slab_bytes = rows_cta * 128;
    const int slab = off / slab_bytes, in = off % slab_bytes;
    const int r = in / 128, chunk_sw = (in % 128) / 16, within = (in % 16) / 2;
    const int chunk = chunk_sw ^ (r & 7);
    const int k = slab * kSlabK + chunk * 8 + within;
    const int n = rank * rows_cta + r;
    const float v = (n < n_valid) ? W[(int64_t)n * ld + k] : 0.f;
    reinterpret_cast<__nv_bfloat16*>(out + (size_t)rank * kWeightBytes)[byte / 2] = __float2bfloat16_rn(v);
  }
  if (blockIdx.x == 0) {
    float* gb = reinterpret_cast<float*>(out + 2 * (size_t)kWeightBytes);
    for (int i = threadIdx.x; i < kBiasFloats; i += blockDim.x)
      gb[i] = i < kN1 ? b1[i] : (i < kN1 + kN2 ? b2[i - kN1] : b3[i - kN1 - kN2]);
  }
}

template <typename TokT, bool kGather>
static int launch(const Params& p, cudaStream_t s) {
  auto kern = fused_mlp_kernel<TokT, kGather>;
  const int smem = kSmemBytes + 1024;
  static bool configured = false;  // per instantiation
  if (!configured) {
    TDSSF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  const int clusters = std::max(1, std::min(num_sms() / 2, p.n_pairs));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TDSSF_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
#ifdef TDSSF_PROFILE
  {  // experiment builds: dump the per-role cycle counters of this launch (sums over warps) and reset them
    static int launches = 0;
    unsigned long long h[64];
    cudaStreamSynchronize(s);
    cudaMemcpy(h, reinterpret_cast<unsigned long long*>(p.debug) + 2, sizeof(h), cudaMemcpyDeviceToHost);
    cudaMemset(p.debug, 0, 1024);
    if (++launches == 8) {
      const double ctas = 2.0 * clusters;
      const char* names[64] = {};
      names[1] = "mma wait own weights"; names[11] = "mma wait peer weights"; names[2] = "mma wait acc_free";
      names[3] = "mma wait a_full"; names[4] = "mma wait h1_full"; names[5] = "mma wait h2_full";
      names[7] = "epi wait acc1_full"; names[8] = "epi wait acc2_full"; names[9] = "epi wait acc3_full";
      names[10] = "prod wait a_empty"; names[20] = "prod consume (incl. a_empty + data wait)"; names[21] = "prod issue";
      names[22] = "prod fill_slots"; names[23] = "prod barrier+compact"; names[24] = "prod total";
      names[25] = "mma total"; names[26] = "epi total"; names[27] = "epi1 body";
      fprintf(stderr, "[tdssf profile] gather=%d tok_bytes=%d n_pairs=%d clusters=%d (cycles per warp, averaged)\n", (int)kGather, (int)sizeof(TokT), p.n_pairs, clusters);
      for (int i = 0; i < 64; ++i)
        if (names[i]) {
          const double warps = (i >= 20 && i <= 24) || i == 10 ? 16.0 * ctas : ((i >= 7 && i <= 9) || i == 26 || i == 27 ? 4.0 * ctas : ctas / 2);
          fprintf(stderr, "  %-44s %12.0f\n", names[i], (double)h[i] / warps);
        }
    }
  }
#endif
  return TDSSF_OK;
}

static unsigned int* debug_word() {
  static unsigned int* d = nullptr;
  if (!d) {
    if (cudaMalloc(&d, 1024) != cudaSuccess) return nullptr;
    cudaMemset(d, 0, 1024);
  }
  return d;
}

}  // namespace tc
}  // namespace tdssf

extern "C" size_t tdssf_mlp_pack_bytes(void) { return tdssf::tc::kPackBytes; }

extern "C" int tdssf_mlp_pack_weights(const tdssf_mlp_weights* w, void* packed, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(w && w->W1 && w->b1 && w->W2 && w->b2 && w->W3 && w->b3 && packed, "tdssf_mlp_pack_weights: NULL pointer");
  TDSSF_REQUIRE(((uintptr_t)packed & 15) == 0, "tdssf_mlp_pack_weights: packed must be 16-byte aligned");
  tc::pack_weights_kernel<<<148, 256, 0, (cudaStream_t)stream>>>(w->W1, w->b1, w->W2, w->b2, w->W3, w->b3, (uint8_t*)packed);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

extern "C" int tdssf_fusion_fwd(const float* xyz, int64_t sweep_stride, int64_t point_stride, const float* K,
                                const float* Rt, const float* img_hw, int B, int V, int NC, int resize_h,
                                int resize_w, int patch, const void* tokens, int tok_dtype, int M, int Ci,
                                const float* x_pt, const void* packed_weights, float* logits, tdssf_rec* rec_out,
                                void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(xyz && K && Rt && img_hw && tokens && x_pt && packed_weights && logits, "tdssf_fusion_fwd: NULL pointer");
  TDSSF_REQUIRE(Ci == tc::kCi, "tdssf_fusion_fwd: built for Ci=%d token channels, got %d", tc::kCi, Ci);
  TDSSF_REQUIRE(NC >= 1 && NC <= TDSSF_MAX_CAMS, "tdssf_fusion_fwd: NC=%d not in 1..%d", NC, TDSSF_MAX_CAMS);
  TDSSF_REQUIRE(B >= 0 && V >= 0 && B < 65536, "tdssf_fusion_fwd: bad B/V");
  TDSSF_REQUIRE(patch > 0 && resize_h >= patch && resize_w >= patch, "tdssf_fusion_fwd: bad resize/patch");
  TDSSF_REQUIRE((resize_h / patch) * (resize_w / patch) == M && M <= 65536, "tdssf_fusion_fwd: M=%d does not match the patch grid", M);
  TDSSF_REQUIRE(point_stride >= 3, "tdssf_fusion_fwd: point_stride < 3");
  TDSSF_REQUIRE((((uintptr_t)tokens | (uintptr_t)x_pt | (uintptr_t)logits | (uintptr_t)packed_weights) & 15) == 0 &&
                    (!rec_out || ((uintptr_t)rec_out & 15) == 0),
                "tdssf_fusion_fwd: pointers must be 16-byte aligned");
  const int64_t N = (int64_t)B * V;
  if (N == 0) return TDSSF_OK;
  if (int rc = tdssf_check_device()) return rc;
  tc::Params p = {};
  p.xyz = xyz; p.sweep_stride = sweep_stride; p.point_stride = point_stride;
  p.K = K; p.Rt = Rt; p.img_hw = img_hw;
  p.V = V; p.NC = NC; p.resize_h = resize_h; p.resize_w = resize_w; p.patch = patch; p.M = M;
  p.tokens = tokens; p.rec_in = nullptr; p.rec_out = rec_out; p.x_img = nullptr; p.x_pt = x_pt;
  p.packed = (const uint8_t*)packed_weights; p.logits = logits; p.N = N;
  p.n_pairs = (int)ceil_div(N, 2 * tc::kTileM);
  p.debug = tc::debug_word();
  cudaStream_t s = (cudaStream_t)stream;
  switch (tok_dtype) {
    case TDSSF_F32: return tc::launch<float, true>(p, s);
    case TDSSF_F16: return tc::launch<__half, true>(p, s);
    case TDSSF_BF16: return tc::launch<__nv_bfloat16, true>(p, s);
    default: set_error("tdssf_fusion_fwd: unknown tok_dtype %d", tok_dtype); return TDSSF_EINVAL;
  }
}

extern "C" int tdssf_mlp_fwd_tc(const float* x_pt, const float* x_img, int64_t N, const void* packed_weights,
                                float* logits, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(x_pt && x_img && packed_weights && logits, "tdssf_mlp_fwd_tc: NULL pointer");
  TDSSF_REQUIRE(N >= 0, "tdssf_mlp_fwd_tc: negative N");
  TDSSF_REQUIRE((((uintptr_t)x_pt | (uintptr_t)x_img | (uintptr_t)logits | (uintptr_t)packed_weights) & 15) == 0,
                "tdssf_mlp_fwd_tc: pointers must be 16-byte aligned");
  if (N == 0) return TDSSF_OK;
  if (int rc = tdssf_check_device()) return rc;
  tc::Params p = {};
  p.x_img = x_img; p.x_pt = x_pt; p.packed = (const uint8_t*)packed_weights; p.logits = logits; p.N = N;
  p.V = 1; p.NC = 1; p.M = 1;
  p.n_pairs = (int)ceil_div(N, 2 * tc::kTileM);
  p.debug = tc::debug_word();
  return tc::launch<float, false>(p, (cudaStream_t)stream);
}
