// a4+a5: nearest-patch token gather and masked mean over cameras (utils/fusion_model.py:184-201).
// One warp per point: the 16-byte record is broadcast-loaded, each valid camera contributes one
// Ci-channel token row read as 16-byte vectors (lane-strided, fully coalesced), summed in camera
// order 0..NC-1 (the order of torch's sum(dim=1)), divided by max(count,1).
// Token maps of a sweep (26-52 MB) stay L2-resident; the output write (1 536 B/pt) is streamed.
#include "common.cuh"

namespace tdssf {

constexpr int kGatherThreads = 256;

template <typename T> struct TokVec;
template <> struct TokVec<float> {
  static constexpr int kPer = 4;  // channels per 16-byte load
  __device__ static void load(const float* p, float (&o)[4]) {
    float4 q = __ldg(reinterpret_cast<const float4*>(p));
    o[0] = q.x; o[1] = q.y; o[2] = q.z; o[3] = q.w;
  }
  __device__ static float round_to_tok(float x) { return x; }
};
template <> struct TokVec<__half> {
  static constexpr int kPer = 8;
  __device__ static void load(const __half* p, float (&o)[8]) {
    uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
    const __half2* h = reinterpret_cast<const __half2*>(&q);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __half22float2(h[i]); o[2 * i] = f.x; o[2 * i + 1] = f.y; }
  }
  __device__ static float round_to_tok(float x) { return __half2float(__float2half_rn(x)); }
};
template <> struct TokVec<__nv_bfloat16> {
  static constexpr int kPer = 8;
  __device__ static void load(const __nv_bfloat16* p, float (&o)[8]) {
    uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); o[2 * i] = f.x; o[2 * i + 1] = f.y; }
  }
  __device__ static float round_to_tok(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
};

// kChunks = Ci / (32 * kPer) when that is exact (compile-time unrolled), 0 = runtime loop.
template <typename T, int kChunks>
__global__ void __launch_bounds__(kGatherThreads)
gather_mean_kernel(const T* __restrict__ tokens, const tdssf_rec* __restrict__ rec,
                   int64_t n_points, int V, int NC, int M, int Ci, float* __restrict__ feat) {
  constexpr int kPer = TokVec<T>::kPer;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int chunks = kChunks > 0 ? kChunks : (Ci + 32 * kPer - 1) / (32 * kPer);

  for (int64_t pt = warp0; pt < n_points; pt += nwarps) {
    const int b = (int)(pt / V);
    uint4 raw = __ldg(reinterpret_cast<const uint4*>(rec) + pt);
    const tdssf_rec r = *reinterpret_cast<const tdssf_rec*>(&raw);
    const T* base = tokens + (int64_t)b * NC * M * Ci;
    const float inv_den = (float)max((int)r.count, 1);

    const int nch = kChunks > 0 ? kChunks : chunks;
#pragma unroll
    for (int ch = 0; ch < nch; ++ch) {
      const int c0 = (ch * 32 + lane) * kPer;
      if (c0 >= Ci) continue;
      float acc[kPer];
#pragma unroll
      for (int i = 0; i < kPer; ++i) acc[i] = 0.f;
#pragma unroll
      for (int cam = 0; cam < TDSSF_MAX_CAMS; ++cam) {
        if (r.mask & (1u << cam)) {  // bits are only ever set for cam < NC
          float t[kPer];
          TokVec<T>::load(base + ((int64_t)cam * M + r.patch[cam]) * Ci + c0, t);
#pragma unroll
          for (int i = 0; i < kPer; ++i) acc[i] = __fadd_rn(acc[i], t[i]);
        }
      }
      float o[kPer];
#pragma unroll
      for (int i = 0; i < kPer; ++i) {
        // same-dtype arithmetic of the reference: sum rounded to the token type, then the quotient
        float s = TokVec<T>::round_to_tok(acc[i]);
        o[i] = TokVec<T>::round_to_tok(__fdiv_rn(s, inv_den));
      }
      float4* dst = reinterpret_cast<float4*>(feat + pt * Ci + c0);
#pragma unroll
      for (int i = 0; i < kPer / 4; ++i)
        st_stream_f4(dst + i, make_float4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]));
    }
  }
}

template <typename T>
static int launch_gather(const void* tokens, const tdssf_rec* rec, int B, int V, int NC, int M, int Ci,
                         float* feat, cudaStream_t s) {
  constexpr int kPer = TokVec<T>::kPer;
  const int64_t n_points = (int64_t)B * V;
  const int warps_per_block = kGatherThreads / 32;
  int64_t blocks = ceil_div(n_points, warps_per_block);
  const int64_t cap = (int64_t)num_sms() * 8 * 4;  // a few waves of resident CTAs, then grid-stride
  if (blocks > cap) blocks = cap;
  const T* tk = reinterpret_cast<const T*>(tokens);
  if (Ci == 384 && kPer == 4)
    gather_mean_kernel<T, 3><<<(unsigned)blocks, kGatherThreads, 0, s>>>(tk, rec, n_points, V, NC, M, Ci, feat);
  else
    gather_mean_kernel<T, 0><<<(unsigned)blocks, kGatherThreads, 0, s>>>(tk, rec, n_points, V, NC, M, Ci, feat);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

}  // namespace tdssf

extern "C" int tdssf_gather_mean(const void* tokens, int tok_dtype, const tdssf_rec* rec, int B, int V,
                                 int NC, int M, int Ci, float* feat, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(tokens && rec && feat, "tdssf_gather_mean: NULL pointer");
  TDSSF_REQUIRE(NC >= 1 && NC <= TDSSF_MAX_CAMS, "tdssf_gather_mean: NC=%d not in 1..%d", NC, TDSSF_MAX_CAMS);
  TDSSF_REQUIRE(M >= 1 && M <= 65536, "tdssf_gather_mean: M=%d not in 1..65536", M);
  TDSSF_REQUIRE(Ci >= 8 && Ci % 8 == 0, "tdssf_gather_mean: Ci=%d must be a positive multiple of 8", Ci);
  TDSSF_REQUIRE(((uintptr_t)tokens & 15) == 0 && ((uintptr_t)feat & 15) == 0 && ((uintptr_t)rec & 15) == 0,
                "tdssf_gather_mean: pointers must be 16-byte aligned");
  if ((int64_t)B * V == 0) return TDSSF_OK;
  cudaStream_t s = (cudaStream_t)stream;
  switch (tok_dtype) {
    case TDSSF_F32: return launch_gather<float>(tokens, rec, B, V, NC, M, Ci, feat, s);
    case TDSSF_F16: return launch_gather<__half>(tokens, rec, B, V, NC, M, Ci, feat, s);
    case TDSSF_BF16: return launch_gather<__nv_bfloat16>(tokens, rec, B, V, NC, M, Ci, feat, s);
    default: set_error("tdssf_gather_mean: unknown tok_dtype %d", tok_dtype); return TDSSF_EINVAL;
  }
}
