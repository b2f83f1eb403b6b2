// a1..a6 fused inference on the 5th-generation tensor cores (tcgen05, bf16 operands, fp32 accumulate in
// TMEM):   project -> mask -> patch gather -> masked camera mean -> concat -> 448->256->64->16 MLP -> logits
// Replaces FeatureFusionModel.forward up to voxel_scores (utils/fusion_model.py:152-204).
//
// One persistent 2-CTA cluster per SM pair (cta_group::2).  Each CTA owns a 128-point tile; the pair
// shares every B operand: W1 (256x448 bf16 = 224 KB) does not fit one SM's shared memory, split in two
// N-halves of 112 KB it stays resident for the whole kernel, and so do W2 and W3.
//
// Per CTA (20 warps, five warpgroups; register budgets rebalanced with setmaxnreg, launch: 96 per thread):
//   prologue     all warps: projection lidar->camera->pixel->patch of every point the CTA will process, one source
//                pointer per point into the workspace (token row / zero row / scratch row holding the camera mean of
//                a multi-camera point), camera records to rec_out.  Done up front because once the gather saturates
//                the load/store unit every memory instruction of another warp queues for hundreds of cycles.
//   warps 0-7    epilogue (72 regs): TMEM -> registers -> ReLU -> bf16 -> TMEM (A operand of the next layer; the biases are
//                already in the accumulators: one ones x bias MMA per layer), in
//                16-column chunks with the next TMEM load in flight; last layer -> logits to HBM
//   warp  8      TMEM allocation, weight load (cp.async.bulk), layer-1 MMA issue (leader CTA only)       (64 regs)
//   warp  9      layer-2 / layer-3 MMA issue as soon as the epilogue has written H1 / H2 (leader CTA only)
//   warps 10-11  idle after the prologue (keep the warpgroups aligned; their registers go to the producers); in training
//                launches one thread of warp 10 stores every finished A slab to x_bf16 (TMA tensor store)
//   warps 12-19  producers (136 regs): per 64-channel K slab, gather of the rows' sources (L2/HBM) with
//                [pointer + immediate] loads, fp32 -> bf16, written in the 128B-swizzled K-major layout the UMMA
//                descriptor expects.  A thread owns a quarter of four rows; three register sets: the loads of slabs
//                s+1 and s+2 are in flight while slab s is converted.
//
// TMEM (512 columns) holds two 256-column tile buffers; inside a buffer the layer-1 accumulator is overwritten in
// place by its own bf16 activation (H1), and the layer-2/3 accumulators and H2 reuse columns already consumed:
//   [0,256) acc1   ->   [0,64) H1 lo | [64,128) acc2 | [128,192) H1 hi | [192,224) H2 | [224,256) acc3
// One thread issues layer 1 of tile t+1 slab by slab as the producers deliver; a second thread issues the small layers of
// tile t (layer 2 when H1 is written, layer 3 when H2 is written), so neither waits behind the other's barriers.
#include <algorithm>
#include <type_traits>
#include "project.cuh"
#include "tc_ptx.cuh"
#include "tmap.cuh"

namespace tdssf {
namespace tc {

constexpr int kTileM = 128;
constexpr int kCp = 64, kCi = 384, kK1 = kCp + kCi;  // 448
constexpr int kN1 = 256, kN2 = 64, kN3 = 32 /* 16 classes padded to 32 */, kClasses = 16;
constexpr int kSlabK = 64;                           // K elements per 128-byte swizzled row
constexpr int kSlabs1 = kK1 / kSlabK;                // 7
constexpr int kSlabs2 = kN1 / kSlabK;                // 4
#ifndef TDSSF_STAGES
#define TDSSF_STAGES 3  // A-operand ring depth; every 16 KB not taken from the unified array is L1 for the gather loads
#endif
constexpr int kStages = TDSSF_STAGES;
constexpr int kEpiWarps = 8, kAuxWarps = 4, kProducerWarps = 8;
constexpr int kMmaWarp = kEpiWarps, kProdWarp0 = kEpiWarps + kAuxWarps;
constexpr int kThreads = 32 * (kEpiWarps + kAuxWarps + kProducerWarps);  // 512
constexpr int kRowsPerWarp = kTileM / kProducerWarps;                    // 16
constexpr int kSteps = kRowsPerWarp / 4;                                 // 4 rows per quarter-warp
// launch: 640 threads x 96 registers = 61440, redistributed over the roles (all of it is used)
// three producer register sets: the loads of two slabs in flight while a third is converted (two sets: 3-5 % slower)
constexpr int kRegsEpi = 72, kRegsAux = 64, kRegsProd = 136;   // 256*72 + 128*64 + 256*136 = 61440

// per-CTA shared-memory image (bytes)
constexpr int kW1Slab = (kN1 / 2) * 128;   // 128 rows x 128 B = 16 KB
constexpr int kW2Slab = (kN2 / 2) * 128;   // 32 rows  -> 4 KB
constexpr int kW3Slab = (kN3 / 2) * 128;   // 16 rows  -> 2 KB
constexpr int kOffW1 = 0;
constexpr int kOffW2 = kOffW1 + kSlabs1 * kW1Slab;          // 114688
constexpr int kOffW3 = kOffW2 + kSlabs2 * kW2Slab;          // 131072
constexpr int kWeightBytes = kOffW3 + kW3Slab;              // 133120 per CTA rank
constexpr int kBiasFloats = kN1 + kN2 + kClasses;           // 336
// The biases ride on the tensor core: every accumulator starts as ones[128 x 16] @ biasrows[N x 16]^T, one extra K = 16
// MMA per layer, instead of 336 shared-memory loads and 208 additions per point row in the epilogues (which queue behind
// the producers' gather in the load/store unit).  K slot 0 carries bf16(b), slot 1 bf16(b - bf16(b)): the sum is b to
// 2^-17.  Both operands are K-major without swizzle: element (row, k) at (row/8)*256 + (k/8)*128 + (row%8)*16 + (k%8)*2.
constexpr int kOffBiasB1 = kWeightBytes;                    // [128 rows][16]  4 KB
constexpr int kOffBiasB2 = kOffBiasB1 + (kN1 / 2) * 32;     // [32 rows][16]   1 KB
constexpr int kOffBiasB3 = kOffBiasB2 + (kN2 / 2) * 32;     // [16 rows][16]   512 B
constexpr int kImageBytes = (kOffBiasB3 + (kN3 / 2) * 32 + 1023) & ~1023;   // 139264 per CTA rank, what the bulk copy brings in
constexpr int kOffRing = kImageBytes;                       // 1024-aligned
constexpr int kASlab = kTileM * 128;                        // 16 KB
constexpr int kOffOnes = kOffRing + kStages * kASlab;       // A operand of the bias MMAs: [128 rows][16], ones in K slots 0 and 1
constexpr int kOffBar = kOffOnes + kTileM * 32;
constexpr int kNumBars = 3 * kStages + 2 + 6 * 2;
constexpr int kOffTmemPtr = kOffBar + kNumBars * 8;
constexpr int kOffGroupCtr = kOffTmemPtr + 8;              // prologue: next unclaimed group of 32 points
constexpr int kSmemBytes = kOffTmemPtr + 16;
constexpr int kCalFloats = 24;                              // per camera: rows 0..2 of lidar->camera, K, 3 floats of padding
constexpr int kCalStageFloats = kStages * kASlab / 4;       // the idle A ring holds the calibration during the prologue
static_assert(kSmemBytes + 1024 <= 232448, "shared memory budget exceeded");
static_assert(kOffRing % 1024 == 0 && kOffW2 % 1024 == 0 && kOffW3 % 1024 == 0 && kImageBytes % 16 == 0, "slab alignment");
static_assert(32 * (kEpiWarps * kRegsEpi + kAuxWarps * kRegsAux + kProducerWarps * kRegsProd) <= (65536 / kThreads / 8 * 8) * kThreads, "register budget");

constexpr size_t kPackBytes = 2 * (size_t)kImageBytes + 64;

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
  kBarALocal = kBarAccFree + 2,        // [kStages] training only: this CTA's 8 producer warps have written the slab (TMA store of x)
};

struct Params {
  // source of the 384 image channels: gather (tokens + projection or records) or a dense matrix
  const float* xyz; int64_t sweep_stride, point_stride;
  const float* K; const float* Rt; const float* img_hw;
  int B, V, NC, resize_h, resize_w, patch, M;
  const void* tokens;
  tdssf_rec* rec_out;          // optional
  uint16_t* x_out;             // optional [N,448] bf16: the MLP input [point feature | image feature] as the tensor cores saw it
  uint16_t* h1_out;            // optional [N,256] bf16: post-ReLU activations of layer 1, saved for the backward pass
  uint16_t* h2_out;            // optional [N,64] bf16
  int64_t* preds_out;          // optional: argmax over the 16 logits + 1 (the class ids of utils/losses.py:66 / utils/test.py)
  const float* x_img;          // dense mode
  const float* x_pt;           // [N,64]
  const uint8_t* packed;       // weights image (tdssf_mlp_pack_weights)
  float* logits;               // [N,16]
  int64_t N;
  int n_pairs;                 // number of 256-row tile pairs
  const void* zeros;           // >= 4 KB of zeros (source of rows without a camera / beyond N)
  const char** src_table;      // [gridDim][iters_max*128] source row of every point this CTA processes (written by its prologue)
  void* scratch;               // [gridDim][iters_max*128] token rows: camera means of multi-camera points
  int iters_max;               // tiles per CTA, upper bound
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
// timeline of one CTA (cluster 5, rank 0): clock64 at event `ev` of tile `it`
#define TRACE(dbg, ev, it)                                                                                        \
  do {                                                                                                              \
    if ((dbg) && (threadIdx.x & 31) == 0 && ptx::cluster_id_x() == 5 && ptx::cluster_ctarank() == 0 && (it) < 32)   \
      reinterpret_cast<unsigned long long*>(dbg)[128 + (ev) * 32 + (it)] = (unsigned long long)clock64();           \
  } while (0)
#else
#define PROF_T0(var)
#define PROF_ADD(dbg, code, var)
#define TRACE(dbg, ev, it)
#endif

#define TDSSF_WAIT_PROBE ptx::mbar_try_wait    // suspends the warp until the phase completes (or the time hint expires)
__device__ __forceinline__ void wait_bar(uint32_t bar, uint32_t parity, unsigned int* debug, int code) {
  if (TDSSF_WAIT_PROBE(bar, parity)) return;
  const long long t0 = clock64();
  while (!TDSSF_WAIT_PROBE(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {  // ~2 s: a protocol bug, never a legitimate wait
      if (debug) atomicExch(debug, (unsigned)code | 0x80000000u);
      __trap();
    }
  }
#ifdef TDSSF_PROFILE
  PROF_ADD(debug, code, t0);
#endif
}

// wait of the single MMA-issuing threads: the suspending try_wait as well (busy polling with test_wait reacts a few tens
// of cycles sooner but keeps the load/store unit busy with barrier reads: measured 0.5 % slower)
__device__ __forceinline__ void wait_bar_poll(uint32_t bar, uint32_t parity, unsigned int* debug, int code) {
  wait_bar(bar, parity, debug, code);
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
__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
// coherent 16-byte load: the scratch rows of multi-camera points are written by this CTA during the kernel
__device__ __forceinline__ uint4 ldg16c(const void* p) {
  uint4 r;
  // L1-allocating: L2-only (.cg) and L1::no_allocate loads were measured 14 % slower
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
struct Raw8 { uint4 lo, hi; };
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
  __device__ static __forceinline__ void load(Raw8& r, const float* piece, int ql) { r = raw8_load_f32(piece, ql); }
  __device__ static __forceinline__ void to_f8(const Raw8& q, float (&f)[8]) { raw8_f32_to_f8(q, f); }
};
template <> struct TokRaw<__half> {
  static constexpr bool kSplit = false;
  __device__ static __forceinline__ void load(Raw8& r, const __half* piece, int ql) { r.lo = ldg16(piece + ql * 8); }
  __device__ static __forceinline__ void to_f8(const Raw8& q, float (&f)[8]) {
    uint32_t w[4] = {q.lo.x, q.lo.y, q.lo.z, q.lo.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 a = __half22float2(*reinterpret_cast<__half2*>(&w[i])); f[2 * i] = a.x; f[2 * i + 1] = a.y; }
  }
};
template <> struct TokRaw<__nv_bfloat16> {
  static constexpr bool kSplit = false;
  __device__ static __forceinline__ void load(Raw8& r, const __nv_bfloat16* piece, int ql) { r.lo = ldg16(piece + ql * 8); }
  __device__ static __forceinline__ void to_f8(const Raw8& q, float (&f)[8]) {
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

// Row of the warp's 16 that quarter-warp q handles in step st.  Within one step the four quarters write rows whose
// swizzle phases (r & 7) are {a, a+4, a+1, a+5}: two of them land in the low 64 bytes of the 128-byte bank window and
// two in the high 64 bytes, so an 8-byte-per-lane store is the minimum two wavefronts.
__device__ __forceinline__ int row_of(int st, int q) {
  return ((q & 1) << 2) | (q >> 1) | ((st & 1) << 1) | ((st >> 1) << 3);
}

// ---- helper-warp routines -------------------------------------------------------------------------------------
// 16 bytes of a token row as floats (4 for fp32, 8 for 16-bit) and back
template <typename TokT> __device__ __forceinline__ void chunk_to_f(const uint4& q, float (&f)[8]) {
  if (sizeof(TokT) == 4) {
    f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y); f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w);
    f[4] = f[5] = f[6] = f[7] = 0.f;
  } else {
    Raw8 r; r.lo = q;
    TokRaw<TokT>::to_f8(r, f);
  }
}
template <typename TokT> __device__ __forceinline__ uint4 f_to_chunk(const float (&a)[8]) {
  if (sizeof(TokT) == 4) {
    return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]), __float_as_uint(a[3]));
  } else if (std::is_same<TokT, __half>::value) {
    __half2 h0 = __floats2half2_rn(a[0], a[1]), h1 = __floats2half2_rn(a[2], a[3]), h2 = __floats2half2_rn(a[4], a[5]),
            h3 = __floats2half2_rn(a[6], a[7]);
    return make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1), *reinterpret_cast<uint32_t*>(&h2),
                      *reinterpret_cast<uint32_t*>(&h3));
  } else {
    return make_uint4(pack_bf16(a[0], a[1]), pack_bf16(a[2], a[3]), pack_bf16(a[4], a[5]), pack_bf16(a[6], a[7]));
  }
}

// Camera mean of two rows that are each seen by exactly two cameras (the common multi-camera case): the whole warp,
// 16 bytes per lane and pass.  Split in two halves so that the caller can put independent work between the loads and
// their use.  (a + b) / 2 in fp32, rounded to the token type for 16-bit tokens (utils/fusion_model.py:198-201).
// second == false: only row 0 is real.
template <typename TokT> struct Mean2 {
  static constexpr int kChunks = kCi * (int)sizeof(TokT) / 16;  // 96 (fp32) or 48 (16-bit)
  static constexpr int kPasses = (kChunks + 31) / 32;
  uint4 qa0[kPasses], qb0[kPasses], qa1[kPasses], qb1[kPasses];
  __device__ __forceinline__ void issue(const char* a0, const char* b0, const char* a1, const char* b1, bool second, int lane) {
#pragma unroll
    for (int i = 0; i < kPasses; ++i) {
      const int c = i * 32 + lane;
      if (c < kChunks) {
        qa0[i] = ldg16(a0 + c * 16);
        qb0[i] = ldg16(b0 + c * 16);
        if (second) {
          qa1[i] = ldg16(a1 + c * 16);
          qb1[i] = ldg16(b1 + c * 16);
        }
      }
    }
  }
  __device__ __forceinline__ void finish(char* d0, char* d1, bool second, int lane) const {
#pragma unroll
    for (int i = 0; i < kPasses; ++i) {
      const int c = i * 32 + lane;
      if (c < kChunks) {
        float x[8], y[8];
        chunk_to_f<TokT>(qa0[i], x);
        chunk_to_f<TokT>(qb0[i], y);
#pragma unroll
        for (int j = 0; j < 8; ++j) x[j] = __fmul_rn(__fadd_rn(x[j], y[j]), 0.5f);  // x/2 == x*0.5 exactly
        *reinterpret_cast<uint4*>(d0 + c * 16) = f_to_chunk<TokT>(x);
        if (second) {
          chunk_to_f<TokT>(qa1[i], x);
          chunk_to_f<TokT>(qb1[i], y);
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] = __fmul_rn(__fadd_rn(x[j], y[j]), 0.5f);
          *reinterpret_cast<uint4*>(d1 + c * 16) = f_to_chunk<TokT>(x);
        }
      }
    }
  }
};

// General case (three or more cameras; no such point exists for a ring of six cameras): camera-order sum, true divide.
template <typename TokT>
__device__ __noinline__ void mean_rows_general(const char* tok, uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3, uint32_t t4,
                                               uint32_t t5, uint32_t cn, char* dst, int lane) {
  const uint32_t t[TDSSF_MAX_CAMS] = {t0, t1, t2, t3, t4, t5};
  constexpr int kRowBytes = kCi * (int)sizeof(TokT);
  constexpr int kChunks = kRowBytes / 16;
  const float den = (float)cn;
  for (int c = lane; c < kChunks; c += 32) {
    float acc[8], f[8];
    chunk_to_f<TokT>(ldg16(tok + (int64_t)t[0] * kRowBytes + c * 16), acc);
    for (uint32_t k = 1; k < cn && k < TDSSF_MAX_CAMS; ++k) {
      uint32_t tk = t[1];
#pragma unroll
      for (int j = 2; j < TDSSF_MAX_CAMS; ++j) if ((uint32_t)j == k) tk = t[j];
      chunk_to_f<TokT>(ldg16(tok + (int64_t)tk * kRowBytes + c * 16), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = __fadd_rn(acc[j], f[j]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = __fdiv_rn(acc[j], den);
    *reinterpret_cast<uint4*>(dst + c * 16) = f_to_chunk<TokT>(acc);
  }
}

// Camera assignment of one point: token rows of the cameras that see it (camera order), their number, and the fields of
// the 16-byte camera record.
struct PointCams {
  uint32_t idx0, idx1, idx2, idx3, idx4, idx5, cnt, mask, pk0, pk1, pk2;
};

// All cameras of one point through the exact path.  Only used for a warp in which some point passes the division-free
// filter for more than two cameras (impossible for a ring of six cameras; kept for arbitrary calibrations).
__device__ __noinline__ PointCams project_point_all(const float* K, const float* Rt, int NC, int M, uint32_t b, float x, float y,
                                                    float z, SweepGeom geo) {
  PointCams r = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  const uint32_t tok_base = b * (uint32_t)(NC * M);
  for (int cam = 0; cam < NC; ++cam) {
    CamCalib cal;
    const float* rt = Rt + ((int64_t)b * NC + cam) * 16;
    const float* kk = K + ((int64_t)b * NC + cam) * 9;
    for (int k = 0; k < 12; ++k) (&cal.r[0][0])[k] = __ldg(rt + k);
    for (int k = 0; k < 9; ++k) (&cal.k[0][0])[k] = __ldg(kk + k);
    const Proj pr = project_one(cal, x, y, z, geo.img_h, geo.img_w);
    const uint32_t f = (uint32_t)patch_index(geo, pr.u, pr.v) & 0xffffu;
    const uint32_t w = f << ((uint32_t)(cam & 1) * 16u);
    if ((cam >> 1) == 0) r.pk0 |= w; else if ((cam >> 1) == 1) r.pk1 |= w; else r.pk2 |= w;
    if (pr.valid) {
      const uint32_t t = tok_base + (uint32_t)cam * (uint32_t)M + f;
      switch (r.cnt) {
        case 0: r.idx0 = t; break;
        case 1: r.idx1 = t; break;
        case 2: r.idx2 = t; break;
        case 3: r.idx3 = t; break;
        case 4: r.idx4 = t; break;
        default: r.idx5 = t; break;
      }
      r.mask |= 1u << cam;
      ++r.cnt;
    }
  }
  return r;
}

// ---- prologue: projection + camera assignment of every point the CTA will process ------------------------------
// Runs before the tile loop with all 20 warps, while the load/store unit is idle (once the producers' gather saturates
// it, every memory instruction of another warp waits hundreds of cycles - measured: in-loop helper warps doing this work
// halved the throughput).  One lane per point.  Pass 1 takes every camera up to the homogeneous image coordinates
// (27 flops) and applies a division-free, conservative visibility filter; the (at most two) cameras a point passes for
// are kept in registers and only those go through pass 2: the IEEE divisions, the exact validity test and the patch
// index.  Bit-identical to running the exact path on all cameras, at about a third of the instructions.
// Result per point: ONE source pointer for the producers - the token row of its camera, a row of zeros (no camera), or,
// for the ~6 % of points seen by several cameras, a scratch row in which the warp writes the camera-order mean
// (utils/fusion_model.py:195-201) - and, on request, the 16-byte camera record.
template <typename TokT, typename RowOf>
__device__ __forceinline__ void project_rows(const Params& p, uint32_t cal_addr, uint32_t ctr_addr, int n_iters, RowOf tile_row0,
                                             int lane) {
  constexpr int kRowBytes = kCi * (int)sizeof(TokT);
  constexpr uint32_t kFull = 0xffffffffu;
  const char* const tok = reinterpret_cast<const char*>(p.tokens);
  const uint32_t n_rows = (uint32_t)p.N, V = (uint32_t)p.V;
  const char** const table = p.src_table + (int64_t)blockIdx.x * p.iters_max * kTileM;
  char* const scratch = reinterpret_cast<char*>(p.scratch) + (int64_t)blockIdx.x * p.iters_max * kTileM * kRowBytes;
  const int n_groups = n_iters * (kTileM / 32);
  const bool want_rec = p.rec_out != nullptr;
  SweepGeom geo = make_geom(1.f, 1.f, p.resize_h, p.resize_w, p.patch);  // grid fields are per launch; sizes per sweep
  const uint32_t f_inv = (uint32_t)patch_index(geo, -1.f, -1.f) & 0xffffu;  // patch of a camera that does not see the point
  // groups of 32 points are claimed dynamically (the camera means make their cost uneven); the next group's point is
  // loaded while the current one is processed
  auto claim = [&]() {
    int g = 0;
    if (lane == 0) g = (int)ptx::atom_add_shared(ctr_addr, 1u);
    return __shfl_sync(kFull, g, 0);
  };
  struct Pt { uint32_t row, b; float x, y, z; };
  auto fetch = [&](int g) {
    Pt t = {0xffffffffu, 0u, 0.f, 0.f, 0.f};
    if (g < n_groups) {
      const int i = g * 32 + lane;
      t.row = (uint32_t)tile_row0(i / kTileM) + (uint32_t)(i % kTileM);
      if (t.row < n_rows) {
        t.b = t.row / V;
        const float* q = p.xyz + (int64_t)t.b * p.sweep_stride + (int64_t)(t.row - t.b * V) * p.point_stride;
        t.x = __ldg(q), t.y = __ldg(q + 1), t.z = __ldg(q + 2);
      }
    }
    return t;
  };
  int grp = claim();
  Pt nxt = fetch(grp);
#pragma unroll 1
  while (grp < n_groups) {
    const int i = grp * 32 + lane;  // index of the point in the CTA's list
    const Pt cur = nxt;
    const int grp_next = claim();
    nxt = fetch(grp_next);
    const uint32_t row = cur.row, b = cur.b;
    const float x = cur.x, y = cur.y, z = cur.z;
    PointCams pc = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    uint32_t nc = 0, ca = 0, cb = 0;
    Homog ha = {0.f, 0.f, 0.f, 1.f}, hb = ha;
    PROF_T0(t_p1);
    if (row < n_rows) {
      const float2 hw = ptx::lds64_ro(cal_addr + (uint32_t)(p.B * p.NC) * (kCalFloats * 4u) + b * 8u);
      geo.img_h = hw.x;
      geo.img_w = hw.y;
      // ---- pass 1: all cameras, no division; calibration from the staged copy, six 16-byte loads per camera ----
      const uint32_t cal_b = cal_addr + b * (uint32_t)p.NC * (kCalFloats * 4u);
#pragma unroll
      for (int cam = 0; cam < TDSSF_MAX_CAMS; ++cam) {
        if (cam < p.NC) {
          const uint32_t a = cal_b + (uint32_t)cam * (kCalFloats * 4u);
          const float4 r0 = ptx::lds128_ro(a), r1 = ptx::lds128_ro(a + 16), r2 = ptx::lds128_ro(a + 32);
          const float4 k0 = ptx::lds128_ro(a + 48), k1 = ptx::lds128_ro(a + 64), k2 = ptx::lds128_ro(a + 80);
          const CamCalib cal = {{{r0.x, r0.y, r0.z, r0.w}, {r1.x, r1.y, r1.z, r1.w}, {r2.x, r2.y, r2.z, r2.w}},
                                {{k0.x, k0.y, k0.z}, {k0.w, k1.x, k1.y}, {k1.z, k1.w, k2.x}}};
          const Homog h = project_homog(cal, x, y, z);
          if (maybe_visible(h, geo.img_h, geo.img_w)) {
            if (nc == 0) { ha = h; ca = cam; } else if (nc == 1) { hb = h; cb = cam; }
            ++nc;
          }
        }
      }
    }
    const bool overflow = __any_sync(kFull, nc > 2);
    PROF_ADD(p.debug, 20, t_p1);
    PROF_T0(t_p2);
    if (overflow) {
      if (row < n_rows) {
        geo.scale_h = __fdiv_rn((float)p.resize_h, fmaxf(geo.img_h, 1e-6f));
        geo.scale_w = __fdiv_rn((float)p.resize_w, fmaxf(geo.img_w, 1e-6f));
        pc = project_point_all(p.K, p.Rt, p.NC, p.M, b, x, y, z, geo);
      }
    } else {
      // ---- pass 2: the candidates, exact ----
      if (nc) {
        geo.scale_h = __fdiv_rn((float)p.resize_h, fmaxf(geo.img_h, 1e-6f));  // utils/fusion_model.py:117-121
        geo.scale_w = __fdiv_rn((float)p.resize_w, fmaxf(geo.img_w, 1e-6f));
      }
      const uint32_t tok_base = b * (uint32_t)(p.NC * p.M);
      pc.pk0 = pc.pk1 = pc.pk2 = f_inv | (f_inv << 16);
#pragma unroll 1
      for (uint32_t s = 0; s < 2; ++s) {
        if (!__any_sync(kFull, nc > s)) break;
        if (nc > s) {
          const Proj pr = project_finish(ha, geo.img_h, geo.img_w);
          const uint32_t f = (uint32_t)patch_index(geo, pr.u, pr.v) & 0xffffu;
          if (want_rec) {
            const uint32_t sh = (ca & 1u) * 16u, keep = ~(0xffffu << sh), w = f << sh;
            if ((ca >> 1) == 0) pc.pk0 = (pc.pk0 & keep) | w;
            else if ((ca >> 1) == 1) pc.pk1 = (pc.pk1 & keep) | w;
            else pc.pk2 = (pc.pk2 & keep) | w;
          }
          if (pr.valid) {
            const uint32_t t = tok_base + ca * (uint32_t)p.M + f;
            if (pc.cnt == 0) pc.idx0 = t; else pc.idx1 = t;
            pc.mask |= 1u << ca;
            ++pc.cnt;
          }
        }
        ha = hb;
        ca = cb;
      }
    }
    if (want_rec && row < n_rows)
      reinterpret_cast<uint4*>(p.rec_out)[row] = make_uint4(pc.pk0, pc.pk1, pc.pk2, pc.mask | (pc.cnt << 8));
    const uint32_t cnt = pc.cnt;
    PROF_ADD(p.debug, 21, t_p2);
    PROF_T0(t_mean);
    table[i] = cnt == 0 ? reinterpret_cast<const char*>(p.zeros)
             : cnt == 1 ? tok + (int64_t)pc.idx0 * kRowBytes
                        : scratch + (int64_t)i * kRowBytes;
    // rows with two cameras: the whole warp averages two rows at a time
    uint32_t two = __ballot_sync(kFull, cnt == 2);
    while (two) {
      const int sa = __ffs(two) - 1;
      two &= two - 1;
      const bool second = two != 0;
      const int sb = second ? __ffs(two) - 1 : sa;
      two &= two - 1;
      const uint32_t a0 = __shfl_sync(kFull, pc.idx0, sa), b0 = __shfl_sync(kFull, pc.idx1, sa);
      const uint32_t a1 = __shfl_sync(kFull, pc.idx0, sb), b1 = __shfl_sync(kFull, pc.idx1, sb);
      Mean2<TokT> m;
      m.issue(tok + (int64_t)a0 * kRowBytes, tok + (int64_t)b0 * kRowBytes, tok + (int64_t)a1 * kRowBytes, tok + (int64_t)b1 * kRowBytes, second, lane);
      m.finish(scratch + (int64_t)(grp * 32 + sa) * kRowBytes, scratch + (int64_t)(grp * 32 + sb) * kRowBytes, second, lane);
    }
    uint32_t many = __ballot_sync(kFull, cnt > 2);
    while (many) {
      const int sa = __ffs(many) - 1;
      many &= many - 1;
      mean_rows_general<TokT>(tok, __shfl_sync(kFull, pc.idx0, sa), __shfl_sync(kFull, pc.idx1, sa),
                              __shfl_sync(kFull, pc.idx2, sa), __shfl_sync(kFull, pc.idx3, sa),
                              __shfl_sync(kFull, pc.idx4, sa), __shfl_sync(kFull, pc.idx5, sa),
                              __shfl_sync(kFull, cnt, sa), scratch + (int64_t)(grp * 32 + sa) * kRowBytes, lane);
    }
    PROF_ADD(p.debug, 22, t_mean);
    grp = grp_next;
  }
}

// kGather: image channels come from the token maps through the records; else from p.x_img
// kSave: training launch, the bf16 input rows / hidden activations also go to p.x_out / p.h1_out / p.h2_out
template <typename TokT, bool kGather, bool kSave>
__global__ void __launch_bounds__(kThreads, 1)
fused_mlp_kernel(const Params p, const __grid_constant__ CUtensorMap tm_x) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the 128B-swizzled operand slabs
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int n_iters = (p.n_pairs - (int)ptx::cluster_id_x() + (int)ptx::num_clusters_x() - 1) / (int)ptx::num_clusters_x();
  auto bar = [&](int i) { return sbase + kOffBar + 8 * i; };
  auto tile_row0 = [&](int it) {
    return (((int64_t)ptx::cluster_id_x() + (int64_t)it * ptx::num_clusters_x()) * 2 + rank) * kTileM;
  };
  volatile uint32_t* s_tmem = reinterpret_cast<volatile uint32_t*>(smem + kOffTmemPtr);

  // ---- one-time setup ------------------------------------------------------------------------------
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(bar(kBarAFull + s), 2 * kProducerWarps);
      ptx::mbar_init(bar(kBarAEmpty + s), (kSave && p.x_out != nullptr) ? 2 : 1);   // + the TMA store of the slab has read it
      ptx::mbar_init(bar(kBarALocal + s), kProducerWarps);
    }
    ptx::mbar_init(bar(kBarWFull), 1);
    ptx::mbar_init(bar(kBarPeerW), 1);
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(bar(kBarAcc1Full + b), 1);
      ptx::mbar_init(bar(kBarAcc2Full + b), 1);
      ptx::mbar_init(bar(kBarAcc3Full + b), 1);
      ptx::mbar_init(bar(kBarH1Full + b), 2 * kEpiWarps);
      ptx::mbar_init(bar(kBarH2Full + b), 2 * kEpiWarps);
      ptx::mbar_init(bar(kBarAccFree + b), 2 * 4);  // the four warps of column half 0
    }
    ptx::fence_mbar_init();
  }
  if (kGather) {  // calibration of every camera of the launch into the A ring, which is idle until the tile loop starts
    const int n_cams = p.B * p.NC;  // host: n_cams * kCalFloats + 2 * B <= kCalStageFloats
    float* s_cal = reinterpret_cast<float*>(smem + kOffRing);
    for (int i = threadIdx.x; i < n_cams * kCalFloats; i += kThreads) {
      const int cam = i / kCalFloats, j = i - cam * kCalFloats;
      s_cal[i] = j < 12 ? __ldg(p.Rt + (int64_t)cam * 16 + j) : j < 21 ? __ldg(p.K + (int64_t)cam * 9 + (j - 12)) : 0.f;
    }
    for (int i = threadIdx.x; i < 2 * p.B; i += kThreads) s_cal[n_cams * kCalFloats + i] = __ldg(p.img_hw + i);
    if (threadIdx.x == 0) *reinterpret_cast<volatile uint32_t*>(smem + kOffGroupCtr) = 0u;
  }
  {  // A operand of the bias MMAs: 1.0 in K slots 0 and 1 of every row (written once, read by the tensor core only)
    uint32_t* ones = reinterpret_cast<uint32_t*>(smem + kOffOnes);
    for (int i = threadIdx.x; i < kTileM * 8; i += kThreads) {
      const int wi = i & 63;                                       // word inside a group of 8 rows (256 bytes)
      ones[i] = (wi < 32 && (wi & 3) == 0) ? 0x3f803f80u : 0u;     // first word of a row's K 0..7 chunk: bf16 {1, 1}
    }
    ptx::fence_proxy_async_smem();
  }
  if (warp == kMmaWarp) {
    ptx::tmem_alloc_2cta(sbase + kOffTmemPtr, 512);
    ptx::tmem_relinquish_2cta();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();  // barriers initialised and TMEM allocated in both CTAs before any remote traffic
  ptx::tc_fence_after();
  const uint32_t tmem = *s_tmem;

  if (warp == kMmaWarp && lane == 0) {  // the weights land while the prologue runs
    const uint8_t* src = p.packed + (size_t)rank * kImageBytes;
    ptx::mbar_expect_tx(bar(kBarWFull), kImageBytes);
    for (int off = 0; off < kImageBytes; off += 16384) {
      const int n = min(16384, kImageBytes - off);
      ptx::bulk_g2s(sbase + kOffW1 + off, src + off, n, bar(kBarWFull));
    }
  }
  if (kGather) {
    if (warp == 0) TRACE(p.debug, 11, 0);
    project_rows<TokT>(p, sbase + kOffRing, sbase + kOffGroupCtr, n_iters, tile_row0, lane);
    PROF_T0(t_sync);
    __syncthreads();
    PROF_ADD(p.debug, 23, t_sync);
    if (warp == 0) TRACE(p.debug, 12, 0);  // the CTA's source table and scratch rows are complete and visible to its producers
  }

  if (warp < kEpiWarps) {
    // =============================== epilogue ======================================================
    // Eight warps: warp w owns TMEM lanes 32*(w & 3) .. +31 (one row per thread) and column half (w >> 2) of every
    // accumulator.  TMEM loads are software-pipelined in 16-column chunks: chunk c+1 is in flight while chunk c gets
    // bias + ReLU + bf16 and is stored back as the A operand of the next layer.
    ptx::setmaxnreg_dec<kRegsEpi>();
    const uint32_t q4 = (uint32_t)warp & 3u, ch = (uint32_t)warp >> 2;
    const uint32_t tlane = tmem + ((q4 * 32u) << 16);
    const int row_in_tile = (int)(q4 * 32u) + lane;
    const uint32_t leader_h1 = ptx::mapa(bar(kBarH1Full), 0), leader_h2 = ptx::mapa(bar(kBarH2Full), 0),
                   leader_free = ptx::mapa(bar(kBarAccFree), 0);
    // ReLU + bf16 of 16 accumulator columns (the bias is already in the accumulator: bias MMA)
    auto act16 = [&](const uint32_t (&v)[16], uint32_t (&h)[8]) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        h[j] = pack_bf16(fmaxf(__uint_as_float(v[2 * j]), 0.f), fmaxf(__uint_as_float(v[2 * j + 1]), 0.f));
    };
    for (int it = 0; it < n_iters; ++it) {
      const uint32_t buf = it & 1, par = (it >> 1) & 1;
      const int64_t row = tile_row0(it) + row_in_tile;
      const uint32_t tb = tlane + buf * 256;
      const bool save_h1 = kSave && p.h1_out != nullptr && row < p.N, save_h2 = kSave && p.h2_out != nullptr && row < p.N;
      // ---- layer 1: acc1 cols [128ch, 128ch+128) -> relu(+b1) -> bf16 H1 cols [128ch, 128ch+64), in place --------
      wait_bar(bar(kBarAcc1Full + buf), par, p.debug, 7);
      ptx::tc_fence_after();
      if (warp == 0) TRACE(p.debug, 4, it);
      {
        uint32_t va[16], vb[16], h[8];
        const uint32_t src = tb + ch * 128;
        ptx::tmem_ld16(src, va);
#pragma unroll
        for (int c = 0; c < 8; c += 2) {
          ptx::tmem_wait_ld();
          ptx::tmem_ld16(src + (c + 1) * 16, vb);
          act16(va, h);
          ptx::tmem_st8(src + c * 8, h);
          if (save_h1) {
            ptx::stg256(p.h1_out + row * kN1 + ch * 128 + c * 16, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);   // 32 bytes per row and instruction
          }
          ptx::tmem_wait_ld();
          if (c + 2 < 8) ptx::tmem_ld16(src + (c + 2) * 16, va);
          act16(vb, h);
          ptx::tmem_st8(src + (c + 1) * 8, h);
          if (save_h1) {
            ptx::stg256(p.h1_out + row * kN1 + ch * 128 + (c + 1) * 16, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);   // 32 bytes per row and instruction
          }
        }
      }
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(leader_h1 + 8 * buf);
      __syncwarp();
      if (warp == 0) TRACE(p.debug, 5, it);
      // ---- layer 2: acc2 cols [64+32ch, +32) -> relu(+b2) -> bf16 H2 cols [192+16ch, +16) -------------------------
      wait_bar(bar(kBarAcc2Full + buf), par, p.debug, 8);
      ptx::tc_fence_after();
      if (warp == 0) TRACE(p.debug, 6, it);
      {
        uint32_t va[16], vb[16], h[8];
        const uint32_t src = tb + 64 + ch * 32, dst = tb + 192 + ch * 16;
        ptx::tmem_ld16(src, va);
        ptx::tmem_ld16(src + 16, vb);
        ptx::tmem_wait_ld();
        act16(va, h);
        ptx::tmem_st8(dst, h);
        if (save_h2) {
          ptx::stg256(p.h2_out + row * kN2 + ch * 32, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);   // 32 bytes per row and instruction
        }
        act16(vb, h);
        ptx::tmem_st8(dst + 8, h);
        if (save_h2) {
          ptx::stg256(p.h2_out + row * kN2 + ch * 32 + 16, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);   // 32 bytes per row and instruction
        }
      }
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(leader_h2 + 8 * buf);
      __syncwarp();
      if (warp == 0) TRACE(p.debug, 7, it);
      // ---- layer 3: acc3 + b3 -> logits (column half 0 only: 16 classes) ---------------------------------------------
      if (ch == 0) {
        wait_bar(bar(kBarAcc3Full + buf), par, p.debug, 9);
        ptx::tc_fence_after();
        if (warp == 0) TRACE(p.debug, 8, it);
        uint32_t v[16];
        ptx::tmem_ld16(tb + 224, v);
        ptx::tmem_wait_ld();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_cluster(leader_free + 8 * buf);
        __syncwarp();
        if (warp == 0) TRACE(p.debug, 9, it);
        if (row < p.N) {
          float4* dst = reinterpret_cast<float4*>(p.logits + row * kClasses);
          float best = -INFINITY;
          int arg = 0;
          bool nan_seen = false;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 o = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                                         __uint_as_float(v[4 * j + 3]));
            const float e[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {  // torch.argmax: first maximum, a NaN counts as the maximum
              const bool is_nan = e[k] != e[k];
              if (!nan_seen && (is_nan || e[k] > best || (j == 0 && k == 0))) { best = e[k]; arg = 4 * j + k; nan_seen = is_nan; }
            }
          }
          // 64 bytes per row as two 256-bit stores (half the store instructions and L1 wavefronts of four 128-bit ones)
          if ((reinterpret_cast<uintptr_t>(p.logits) & 31) == 0) {
            ptx::stg256(dst, v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
            ptx::stg256(dst + 2, v[8], v[9], v[10], v[11], v[12], v[13], v[14], v[15]);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) dst[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
          }
          if (p.preds_out) p.preds_out[row] = arg + 1;
        }
      }
    }
  } else if (warp < kProdWarp0) {
    if (kRegsAux < 96) ptx::setmaxnreg_dec<kRegsAux>(); else ptx::setmaxnreg_inc<kRegsAux>();
    constexpr uint32_t kI1 = ptx::idesc_bf16(256, kN1), kI2 = ptx::idesc_bf16(256, kN2), kI3 = ptx::idesc_bf16(256, kN3);
    const uint64_t ones_desc = ptx::smem_desc_nosw(sbase + kOffOnes, 128, 256);
    if (warp == kMmaWarp) {
      // =============================== weights + layer-1 MMA issue ===================================
      // tcgen05.mma issue is paced by the tensor pipe (the issuing thread blocks while the previous MMA starts), so the
      // pipe is busy only while this thread has the next MMA at hand: its loop is kept to wait - 4 MMAs - commit, and
      // the small layers of the previous tile are issued by another thread (warp kMmaWarp + 1).
      if (leader && lane == 0 && n_iters > 0) {
        // B operands are read from both CTAs' shared memory: wait for our image and for the peer's
        wait_bar(bar(kBarWFull), 0, p.debug, 1);
        wait_bar(bar(kBarPeerW), 0, p.debug, 11);
        uint32_t stage = 0, phase = 0;  // A ring cursor
        for (int it = 0; it < n_iters; ++it) {
          const uint32_t buf = it & 1;
          if (it >= 2) wait_bar_poll(bar(kBarAccFree + buf), ((it >> 1) - 1) & 1, p.debug, 2);
          ptx::tc_fence_after();
          TRACE(p.debug, 0, it);
          ptx::umma_ss_2cta(tmem + buf * 256, ones_desc, ptx::smem_desc_nosw(sbase + kOffBiasB1, 128, 256), kI1, 0);  // acc1 = b1
#pragma unroll 1
          for (int s = 0; s < kSlabs1; ++s) {
            wait_bar_poll(bar(kBarAFull + stage), phase, p.debug, 3);
            ptx::tc_fence_after();
            if (it >= 4 && it < 8) TRACE(p.debug, 11, 1 + (it - 4) * 7 + s);   // slab trace: slab seen full
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint64_t ad = ptx::smem_desc_sw128(sbase + kOffRing + stage * kASlab + kk * 32);
              const uint64_t bd = ptx::smem_desc_sw128(sbase + kOffW1 + s * kW1Slab + kk * 32);
              ptx::umma_ss_2cta(tmem + buf * 256, ad, bd, kI1, 1);
            }
            ptx::umma_commit_2cta(bar(kBarAEmpty + stage), 3);
            if (it >= 4 && it < 8) TRACE(p.debug, 12, 1 + (it - 4) * 7 + s);   // its MMAs and commit issued
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
          ptx::umma_commit_2cta(bar(kBarAcc1Full + buf), 3);
          TRACE(p.debug, 1, it);
        }
      }
      if (!leader && lane == 0) {
        wait_bar(bar(kBarWFull), 0, p.debug, 6);
        ptx::mbar_arrive_cluster(ptx::mapa(bar(kBarPeerW), 0));
      }
      __syncwarp();
    } else if (warp == kMmaWarp + 1) {
      // =============================== layer-2 / layer-3 MMA issue ===================================
      // A operands are the bf16 activations the epilogue wrote into TMEM; issued as soon as they are complete, from a
      // thread of its own so that they never wait behind (or delay) the layer-1 stream of the next tile.
      if (leader && lane == 0 && n_iters > 0) {
        wait_bar(bar(kBarWFull), 0, p.debug, 1);
        wait_bar(bar(kBarPeerW), 0, p.debug, 11);
        for (int it = 0; it < n_iters; ++it) {
          const uint32_t buf = it & 1, par = (it >> 1) & 1, tb = tmem + buf * 256;
          wait_bar_poll(bar(kBarH1Full + buf), par, p.debug, 4);
          ptx::tc_fence_after();
          TRACE(p.debug, 2, it);
          // layer 2: A = H1 (TMEM cols [0,64) and [128,192) of the buffer), B = W2, D = cols [64,128), starting from b2
          ptx::umma_ss_2cta(tb + 64, ones_desc, ptx::smem_desc_nosw(sbase + kOffBiasB2, 128, 256), kI2, 0);
#pragma unroll
          for (int k = 0; k < kN1 / 16; ++k) {
            const uint64_t bd = ptx::smem_desc_sw128(sbase + kOffW2 + (k >> 2) * kW2Slab + (k & 3) * 32);
            ptx::umma_ts_2cta(tb + 64, tb + (k < 8 ? k * 8 : 128 + (k - 8) * 8), bd, kI2, 1);
          }
          ptx::umma_commit_2cta(bar(kBarAcc2Full + buf), 3);
          wait_bar_poll(bar(kBarH2Full + buf), par, p.debug, 5);
          ptx::tc_fence_after();
          TRACE(p.debug, 3, it);
          // layer 3: A = H2 (cols [192,224)), B = W3, D = cols [224,256), starting from b3
          ptx::umma_ss_2cta(tb + 224, ones_desc, ptx::smem_desc_nosw(sbase + kOffBiasB3, 128, 256), kI3, 0);
#pragma unroll
          for (int k = 0; k < kN2 / 16; ++k) {
            const uint64_t bd = ptx::smem_desc_sw128(sbase + kOffW3 + k * 32);
            ptx::umma_ts_2cta(tb + 224, tb + 192 + k * 8, bd, kI3, 1);
          }
          ptx::umma_commit_2cta(bar(kBarAcc3Full + buf), 3);
        }
      }
      __syncwarp();
    } else if (kSave && warp == kMmaWarp + 2) {
      // =============================== training: x rows to HBM by TMA ================================
      // The finished A slab IS the bf16 [128 x 64] block of x_out (128B-swizzled, which the tensor map undoes): one bulk
      // tensor store per slab instead of 16 store instructions per producer thread in the load/store unit the gather
      // lives on.  The slab is handed back (second arrival on its empty barrier) once the store has read it.
      if (p.x_out != nullptr && lane == 0) {
        uint32_t stage = 0, phase = 0;
        for (int it = 0; it < n_iters; ++it) {
          const int row0 = (int)tile_row0(it);
#pragma unroll 1
          for (int s = 0; s < kSlabs1; ++s) {
            wait_bar(bar(kBarALocal + stage), phase, p.debug, 12);
            ptx::tma_store_2d(&tm_x, s * kSlabK, row0, sbase + kOffRing + stage * kASlab);
            ptx::bulk_commit_group();
            ptx::bulk_wait_group_read<0>();
            ptx::mbar_arrive_local(bar(kBarAEmpty + stage));
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
        ptx::bulk_wait_group<0>();
      }
      __syncwarp();
    }
  } else {
    // =============================== producers =====================================================
    // The instruction stream per slab is kept minimal (about 50 per thread): one source pointer per row is fetched
    // per tile, the seven slabs are unrolled so that every load is [pointer + immediate], and the stores use
    // precomputed 32-bit shared-memory offsets.  Three register sets: the loads of slabs s+1 and s+2 are in flight
    // while slab s is converted (up to 96 KB per SM).
    ptx::setmaxnreg_inc<kRegsProd>();
    const int pw = warp - kProdWarp0;  // 0..7
    const int quarter = lane >> 3, ql = lane & 7;
    const uint32_t leader_full0 = ptx::mapa(bar(kBarAFull), 0);
    const char* const* const table = p.src_table + (int64_t)blockIdx.x * p.iters_max * kTileM + pw * kRowsPerWarp;
    constexpr int kPieceBytes = kSlabK * (int)sizeof(TokT);   // 64 channels of a token row
    constexpr bool kTok32 = sizeof(TokT) == 4 || !kGather;    // image channels arrive as fp32
    const char* const zero_row = reinterpret_cast<const char*>(p.zeros);

    // shared-memory byte offsets (from the slab start) of the thread's 8-byte (fp32 sources: two per row) or 16-byte
    // stores, for the rows of step 0 and step 1; steps 2 and 3 are 8 rows = 1024 bytes further
    uint32_t so_lo[2], so_hi[2], so_16[2];
#pragma unroll
    for (int st = 0; st < 2; ++st) {
      const uint32_t r = (uint32_t)(pw * kRowsPerWarp + row_of(st, quarter)), sw = r & 7u;
      const uint32_t c0 = (uint32_t)ql >> 1, hb = ((uint32_t)ql & 1u) * 8u;
      so_lo[st] = sbase + kOffRing + r * 128u + ((c0 ^ sw) << 4) + hb;
      so_hi[st] = sbase + kOffRing + r * 128u + (((c0 + 4u) ^ sw) << 4) + hb;
      so_16[st] = sbase + kOffRing + r * 128u + (((uint32_t)ql ^ sw) << 4);
    }

    const char* pa[kSteps];   // image-channel source of the thread's four rows, lane offset included
    const char* pn[kSteps];   // the same for the next tile (read from the CTA's source table one tile ahead)
#pragma unroll
    for (int st = 0; st < kSteps; ++st) { pa[st] = zero_row; pn[st] = zero_row; }
    auto load_sources = [&](int it) {
      if (!kGather || it >= n_iters) return;
#pragma unroll
      for (int st = 0; st < kSteps; ++st) {
        uint64_t v;
        asm volatile("ld.global.b64 %0, [%1];" : "=l"(v) : "l"(table + it * kTileM + row_of(st, quarter)));
        pn[st] = reinterpret_cast<const char*>(v);
      }
    };
    load_sources(0);

    struct Set { Raw8 a[kSteps]; };

    // loads of slab S of tile `it`: setup once per slab, then one step (= one row of the quarter-warp) at a time
    auto issue_setup = [&](auto S_, int it) {
      constexpr int S = decltype(S_)::value;
      if (S == 1) {  // the tile's row sources
        if (kGather) {
#pragma unroll
          for (int st = 0; st < kSteps; ++st) pa[st] = pn[st] + ql * 16;
          load_sources(it + 1);
        } else {
          const int64_t row0 = tile_row0(it) + pw * kRowsPerWarp;
#pragma unroll
          for (int st = 0; st < kSteps; ++st) {
            const int64_t row = row0 + row_of(st, quarter);
            pa[st] = row < p.N ? reinterpret_cast<const char*>(p.x_img) + row * (kCi * 4) + ql * 16 : zero_row;
          }
        }
      }
    };
    auto issue_one = [&](auto S_, int it, int st, Set& n) {
      constexpr int S = decltype(S_)::value;
      if (S == 0) {  // point features
        const int64_t row = tile_row0(it) + pw * kRowsPerWarp + row_of(st, quarter);
        const char* px = row < p.N ? reinterpret_cast<const char*>(p.x_pt) + row * (kCp * 4) + ql * 16 : zero_row;
        n.a[st].lo = ldg16c(px);
        n.a[st].hi = ldg16c(px + 128);
        return;
      }
      constexpr int kOff = (S - 1) * (kTok32 ? 256 : kPieceBytes);
      n.a[st].lo = ldg16c(pa[st] + kOff);
      if (kTok32) n.a[st].hi = ldg16c(pa[st] + kOff + 128);
    };
    auto issue = [&](auto S_, int it, Set& n) {
      issue_setup(S_, it);
#pragma unroll
      for (int st = 0; st < kSteps; ++st) issue_one(S_, it, st, n);
    };

    uint32_t stage = 0, phase = 0;  // A ring cursor
    // conversion + store of one step of slab S out of set c into ring slot `stage`
    auto consume_one = [&](auto S_, int it, int st, const Set& c, uint32_t slab_off) {
      constexpr int S = decltype(S_)::value;
      constexpr bool kF32 = S == 0 || kTok32;  // source elements are fp32
      float acc[8];
      (void)it;   // (training: the finished slab goes to x_out by TMA, warp kMmaWarp + 2)
      if (kF32) {
        raw8_f32_to_f8(c.a[st], acc);
        const uint32_t w0 = pack_bf16(acc[0], acc[1]), w1 = pack_bf16(acc[2], acc[3]), w2 = pack_bf16(acc[4], acc[5]),
                       w3 = pack_bf16(acc[6], acc[7]);
        ptx::sts64(so_lo[st & 1] + slab_off + (st >> 1) * 1024, w0, w1);
        ptx::sts64(so_hi[st & 1] + slab_off + (st >> 1) * 1024, w2, w3);
      } else {
        TokRaw<TokT>::to_f8(c.a[st], acc);
        const uint32_t w0 = pack_bf16(acc[0], acc[1]), w1 = pack_bf16(acc[2], acc[3]), w2 = pack_bf16(acc[4], acc[5]),
                       w3 = pack_bf16(acc[6], acc[7]);
        ptx::sts128(so_16[st & 1] + slab_off + (st >> 1) * 1024, w0, w1, w2, w3);
      }
    };
    // One pipeline step: the loads of slab S2 of tile it2 go into set `nxt` (all eight of the thread at once, before
    // anything can block), then slab S of tile `it` leaves set `cur` for the ring.  (Measured: alternating loads and
    // stores row by row, or issuing the loads after the wait for the ring slot, is 2-5 % slower.)
    auto step_fused = [&](auto S_, int it, const Set& cur, auto S2_, int it2, Set& nxt) {
      constexpr int S = decltype(S_)::value;
      if (it2 < n_iters) issue(S2_, it2, nxt);
      if (S == 0 && pw == 0) TRACE(p.debug, 10, it);
      if (pw == 0 && it >= 4 && it < 8) TRACE(p.debug, 13, (it - 4) * 7 + S);   // slab trace: loads issued
      wait_bar(bar(kBarAEmpty + stage), phase ^ 1, p.debug, 10);
      if (pw == 0 && it >= 4 && it < 8) TRACE(p.debug, 14, (it - 4) * 7 + S);   // ring slot free
      const uint32_t slab_off = stage * kASlab;
#pragma unroll
      for (int st = 0; st < kSteps; ++st) consume_one(S_, it, st, cur, slab_off);
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        ptx::mbar_arrive_cluster(leader_full0 + 8 * stage);
        if (kSave && p.x_out != nullptr) ptx::mbar_arrive_local(bar(kBarALocal + stage));
      }
      __syncwarp();
      if (pw == 0 && it >= 4 && it < 8) TRACE(p.debug, 15, (it - 4) * 7 + S);   // slab written and announced
      if (++stage == kStages) { stage = 0; phase ^= 1; }
    };

    // Item stream (tile, slab): item i is converted out of register set i % kSets after the loads of item
    // i + kSets - 1 went into the set that item i - 1 has just left.  7 slabs per tile, so the rotation closes after
    // kSets tiles.
    Set R0, R1, R2;
    auto step = [&](auto S_, int it, Set& cur, Set& nxt2) {
      constexpr int S = decltype(S_)::value;
      constexpr int S2 = (S + 2) % kSlabs1;
      const int it2 = it + (S + 2 >= kSlabs1 ? 1 : 0);
      step_fused(S_, it, cur, std::integral_constant<int, S2>{}, it2, nxt2);
    };
    auto tile = [&](int it, Set& x, Set& y, Set& z) {  // slab 0 in x, slab 1 in y
      step(std::integral_constant<int, 0>{}, it, x, z);
      step(std::integral_constant<int, 1>{}, it, y, x);
      step(std::integral_constant<int, 2>{}, it, z, y);
      step(std::integral_constant<int, 3>{}, it, x, z);
      step(std::integral_constant<int, 4>{}, it, y, x);
      step(std::integral_constant<int, 5>{}, it, z, y);   // loads slab 0 of the next tile into y
      step(std::integral_constant<int, 6>{}, it, x, z);   // loads slab 1 of the next tile into z
    };
    if (n_iters > 0) {
      issue(std::integral_constant<int, 0>{}, 0, R0);
      issue(std::integral_constant<int, 1>{}, 0, R1);
#pragma unroll 1
      for (int it = 0; it < n_iters; it += 3) {
        tile(it, R0, R1, R2);
        if (it + 1 < n_iters) tile(it + 1, R1, R2, R0);
        if (it + 2 < n_iters) tile(it + 2, R2, R0, R1);
      }
    }
  }

  // ---- teardown ----------------------------------------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == kMmaWarp) ptx::tmem_dealloc_2cta(tmem, 512);
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
    reinterpret_cast<__nv_bfloat16*>(out + (size_t)rank * kImageBytes)[byte / 2] = __float2bfloat16_rn(v);
  }
  // B operands of the bias MMAs: per layer [rows of this rank][16], K slot 0 = bf16(b), slot 1 = bf16(b - slot 0), rest 0
  const int bias_elems = (kImageBytes - kWeightBytes) / 2;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < 2 * bias_elems; e += gridDim.x * blockDim.x) {
    const int rank = e / bias_elems;
    const int byte = kWeightBytes + (e % bias_elems) * 2;
    const float* b; int rows_cta, off, n_valid;
    if (byte < kOffBiasB2) { b = b1; rows_cta = kN1 / 2; off = byte - kOffBiasB1; n_valid = kN1; }
    else if (byte < kOffBiasB3) { b = b2; rows_cta = kN2 / 2; off = byte - kOffBiasB2; n_valid = kN2; }
    else { b = b3; rows_cta = kN3 / 2; off = byte - kOffBiasB3; n_valid = kClasses; }
    const int r = (off / 256) * 8 + (off % 128) / 16, k = ((off % 256) / 128) * 8 + (off % 16) / 2;
    const int n = rank * rows_cta + r;
    float v = 0.f;
    if (r < rows_cta && n < n_valid && k < 2) {
      const float hi = __bfloat162float(__float2bfloat16_rn(b[n]));
      v = k == 0 ? hi : b[n] - hi;
    }
    reinterpret_cast<__nv_bfloat16*>(out + (size_t)rank * kImageBytes)[byte / 2] = __float2bfloat16_rn(v);
  }
}

template <typename TokT, bool kGather, bool kSave>
static int launch_as(const Params& p, const CUtensorMap& tm_x, cudaStream_t s) {
  auto kern = fused_mlp_kernel<TokT, kGather, kSave>;
  const int smem = kSmemBytes + 1024;
  static bool configured[kMaxDevices] = {};  // per instantiation and device (the attribute is per device)
  const int dev = current_device();
  if (!configured[dev]) {
    TDSSF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured[dev] = true;
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
  TDSSF_CUDA(cudaLaunchKernelEx(&cfg, kern, p, tm_x));
#ifdef TDSSF_PROFILE
#include "exp/mlp_tc_profile_dump.inc"
#endif
  return TDSSF_OK;
}

// the training instantiation only when something is to be saved: the inference kernel carries none of that code
template <typename TokT, bool kGather>
static int launch(Params p, cudaStream_t s) {
  CUtensorMap tm_x = {};   // training launches store their bf16 input rows [N, 448] with TMA out of the A ring
  if (p.x_out && !tmap::make_map(&tm_x, p.x_out, p.N, kK1, kTileM)) {
    set_error("tdssf_fusion_fwd: cuTensorMapEncodeTiled failed for x_bf16 (driver entry point unavailable or bad geometry)");
    return TDSSF_ENOSUP;
  }
  return (p.x_out || p.h1_out || p.h2_out) ? launch_as<TokT, kGather, true>(p, tm_x, s) : launch_as<TokT, kGather, false>(p, tm_x, s);
}

__device__ __align__(16) unsigned char g_zero_rows[4096];  // zero-initialised at module load, never written

static const void* zero_rows() {  // a __device__ symbol has one instance per device
  static void* z[kMaxDevices] = {};
  const int dev = current_device();
  if (!z[dev] && cudaGetSymbolAddress(&z[dev], g_zero_rows) != cudaSuccess) z[dev] = nullptr;
  return z[dev];
}

// tiles per CTA (upper bound) for N points on this device
static int iters_max_for(int64_t N) {
  const int n_pairs = (int)ceil_div(N, 2 * kTileM);
  const int clusters = std::max(1, std::min(num_sms() / 2, n_pairs));
  return (int)ceil_div(n_pairs, clusters);
}
// workspace = source table [ctas][iters_max*128] x 8 B, then scratch rows [ctas][iters_max*128] x one token row
static size_t table_bytes(int64_t N) { return ((size_t)num_sms() * iters_max_for(N) * kTileM * 8 + 255) & ~(size_t)255; }
static size_t scratch_bytes(int64_t N, int tok_dtype) {
  return table_bytes(N) + (size_t)num_sms() * iters_max_for(N) * kTileM * kCi * (tok_dtype == TDSSF_F32 ? 4 : 2);
}

__device__ unsigned int g_debug_words[2048];  // [0]: code of a barrier wait that timed out (protocol bug); rest: profile builds

static unsigned int* debug_word() {
  static void* d[kMaxDevices] = {};
  const int dev = current_device();
  if (!d[dev] && cudaGetSymbolAddress(&d[dev], g_debug_words) != cudaSuccess) d[dev] = nullptr;
  return reinterpret_cast<unsigned int*>(d[dev]);
}

}  // namespace tc
}  // namespace tdssf

extern "C" size_t tdssf_mlp_pack_bytes(void) { return tdssf::tc::kPackBytes; }

extern "C" size_t tdssf_fusion_workspace_bytes(int B, int V, int tok_dtype) {
  return tdssf::tc::scratch_bytes((int64_t)std::max(B, 0) * std::max(V, 0), tok_dtype);
}

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
                                int64_t* preds_out, void* x_bf16, void* h1_bf16, void* h2_bf16,
                                void* workspace, size_t workspace_bytes, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(xyz && K && Rt && img_hw && tokens && x_pt && packed_weights && logits, "tdssf_fusion_fwd: NULL pointer");
  TDSSF_REQUIRE(Ci == tc::kCi, "tdssf_fusion_fwd: built for Ci=%d token channels, got %d", tc::kCi, Ci);
  TDSSF_REQUIRE(NC >= 1 && NC <= TDSSF_MAX_CAMS, "tdssf_fusion_fwd: NC=%d not in 1..%d", NC, TDSSF_MAX_CAMS);
  TDSSF_REQUIRE(B >= 0 && V >= 0 && B < 65536, "tdssf_fusion_fwd: bad B/V");
  TDSSF_REQUIRE(patch > 0 && resize_h >= patch && resize_w >= patch, "tdssf_fusion_fwd: bad resize/patch");
  TDSSF_REQUIRE((resize_h / patch) * (resize_w / patch) == M && M <= 32768, "tdssf_fusion_fwd: M=%d does not match the patch grid (or exceeds 32768)", M);
  TDSSF_REQUIRE(point_stride >= 3, "tdssf_fusion_fwd: point_stride < 3");
  if ((int64_t)B * NC * M >= (1ll << 32) || (int64_t)B * V >= (1ll << 31)) {  // 32-bit row arithmetic inside the kernel
    set_error("tdssf_fusion_fwd: B*NC*M = %lld token rows / B*V = %lld points exceed the 32-bit index range",
              (long long)B * NC * M, (long long)B * V);
    return TDSSF_ENOSUP;
  }
  TDSSF_REQUIRE((((uintptr_t)tokens | (uintptr_t)x_pt | (uintptr_t)logits | (uintptr_t)packed_weights) & 15) == 0 &&
                    (!rec_out || ((uintptr_t)rec_out & 15) == 0) &&
                    ((uintptr_t)x_bf16 & 15) == 0,
                "tdssf_fusion_fwd: pointers must be 16-byte aligned");
  TDSSF_REQUIRE((((uintptr_t)h1_bf16 | (uintptr_t)h2_bf16) & 31) == 0, "tdssf_fusion_fwd: h1_bf16 / h2_bf16 must be 32-byte aligned (256-bit stores)");
  const int64_t N = (int64_t)B * V;
  if (N == 0) return TDSSF_OK;
  if (int rc = tdssf_check_device()) return rc;
  TDSSF_REQUIRE(tok_dtype == TDSSF_F32 || tok_dtype == TDSSF_F16 || tok_dtype == TDSSF_BF16, "tdssf_fusion_fwd: unknown tok_dtype %d", tok_dtype);
  TDSSF_REQUIRE(workspace && ((uintptr_t)workspace & 255) == 0 && workspace_bytes >= tc::scratch_bytes(N, tok_dtype),
                "tdssf_fusion_fwd: 256-byte aligned workspace of %zu bytes needed (tdssf_fusion_workspace_bytes), got %zu",
                tc::scratch_bytes(N, tok_dtype), workspace_bytes);
  const void* zeros = tc::zero_rows();
  TDSSF_REQUIRE(zeros != nullptr, "tdssf: could not allocate the zero row");
  cudaStream_t s = (cudaStream_t)stream;
  // The kernel stages the calibration of all its sweeps in shared memory: at most `b_max` sweeps per launch (84 for six
  // cameras), more are processed in consecutive launches that reuse the workspace.
  const int b_max = std::max(1, tc::kCalStageFloats / (NC * tc::kCalFloats + 2));
  const size_t tok_row = (size_t)tc::kCi * (tok_dtype == TDSSF_F32 ? 4 : 2);
  for (int b0 = 0; b0 < B; b0 += b_max) {
    const int Bc = std::min(b_max, B - b0);
    const int64_t Nc = (int64_t)Bc * V, row0 = (int64_t)b0 * V;
    tc::Params p = {};
    p.src_table = reinterpret_cast<const char**>(workspace);
    p.scratch = reinterpret_cast<char*>(workspace) + tc::table_bytes(Nc);
    p.iters_max = tc::iters_max_for(Nc);
    p.xyz = xyz + (int64_t)b0 * sweep_stride; p.sweep_stride = sweep_stride; p.point_stride = point_stride;
    p.K = K + (int64_t)b0 * NC * 9; p.Rt = Rt + (int64_t)b0 * NC * 16; p.img_hw = img_hw + 2 * (int64_t)b0;
    p.B = Bc; p.V = V; p.NC = NC; p.resize_h = resize_h; p.resize_w = resize_w; p.patch = patch; p.M = M;
    p.tokens = reinterpret_cast<const char*>(tokens) + (size_t)b0 * NC * M * tok_row;
    p.rec_out = rec_out ? rec_out + row0 : nullptr; p.preds_out = preds_out ? preds_out + row0 : nullptr; p.x_img = nullptr;
    p.x_out = x_bf16 ? reinterpret_cast<uint16_t*>(x_bf16) + row0 * tc::kK1 : nullptr;
    p.h1_out = h1_bf16 ? reinterpret_cast<uint16_t*>(h1_bf16) + row0 * tc::kN1 : nullptr;
    p.h2_out = h2_bf16 ? reinterpret_cast<uint16_t*>(h2_bf16) + row0 * tc::kN2 : nullptr; p.x_pt = x_pt + row0 * tc::kCp;
    p.packed = (const uint8_t*)packed_weights; p.logits = logits + row0 * tc::kClasses; p.N = Nc;
    p.n_pairs = (int)ceil_div(Nc, 2 * tc::kTileM);
    p.debug = tc::debug_word();
    p.zeros = zeros;
    int rc;
    switch (tok_dtype) {
      case TDSSF_F32: rc = tc::launch<float, true>(p, s); break;
      case TDSSF_F16: rc = tc::launch<__half, true>(p, s); break;
      default: rc = tc::launch<__nv_bfloat16, true>(p, s); break;
    }
    if (rc != TDSSF_OK) return rc;
  }
  return TDSSF_OK;
}

extern "C" int tdssf_mlp_fwd_tc(const float* x_pt, const float* x_img, int64_t N, const void* packed_weights,
                                float* logits, void* h1_bf16, void* h2_bf16, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(x_pt && x_img && packed_weights && logits, "tdssf_mlp_fwd_tc: NULL pointer");
  TDSSF_REQUIRE(N >= 0, "tdssf_mlp_fwd_tc: negative N");
  TDSSF_REQUIRE((((uintptr_t)x_pt | (uintptr_t)x_img | (uintptr_t)logits | (uintptr_t)packed_weights | (uintptr_t)h1_bf16 |
                  (uintptr_t)h2_bf16) & 15) == 0,
                "tdssf_mlp_fwd_tc: pointers must be 16-byte aligned");
  TDSSF_REQUIRE((((uintptr_t)h1_bf16 | (uintptr_t)h2_bf16) & 31) == 0, "tdssf_mlp_fwd_tc: h1_bf16 / h2_bf16 must be 32-byte aligned (256-bit stores)");
  if (N == 0) return TDSSF_OK;
  if (int rc = tdssf_check_device()) return rc;
  tc::Params p = {};
  p.x_img = x_img; p.x_pt = x_pt; p.packed = (const uint8_t*)packed_weights; p.logits = logits; p.N = N;
  p.h1_out = reinterpret_cast<uint16_t*>(h1_bf16); p.h2_out = reinterpret_cast<uint16_t*>(h2_bf16);
  p.B = 1; p.V = 1; p.NC = 1; p.M = 1;
  p.n_pairs = (int)ceil_div(N, 2 * tc::kTileM);
  p.debug = tc::debug_word();
  p.zeros = tc::zero_rows();
  TDSSF_REQUIRE(p.zeros != nullptr, "tdssf: could not allocate the zero row");
  return tc::launch<float, false>(p, (cudaStream_t)stream);
}
