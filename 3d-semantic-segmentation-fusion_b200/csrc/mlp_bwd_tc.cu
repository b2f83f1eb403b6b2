// a6 backward in bf16 (training with the tensor-core forward): the gradients of the 448->256->64->16 MLP from the bf16
// activations tdssf_mlp_fwd_tc saved.  Replaces the autograd of FeatureFusionModel.mlp (utils/fusion_model.py:144-150,
// 203-204; utils/train.py:91 calls backward on the total loss).
//
// The six matrix products are plain GEMMs (no fusion with a gather, no special layout), so they go to cuBLAS
// (cublasGemmEx, bf16 operands, fp32 accumulate, fp32 outputs accumulated in place); hand-written here are the
// element-wise kernels around them: fp32 -> bf16 conversion with the [x_pt | x_img] concatenation, the ReLU masks with
// the bias-gradient column sums, and the orchestration.  cuBLAS is loaded at first use (dlopen), so libtdssf.so itself
// has no link-time dependency on it.
//
//   dl   = bf16(dlogits)                         db3 += sum_rows dlogits
//   G2   = dl  @ W3            [N,64]            dh2 = G2 * (h2 > 0)     db2 += sum_rows dh2
//   G1   = dh2 @ W2            [N,256]           dh1 = G1 * (h1 > 0)     db1 += sum_rows dh1
//   dx_pt = dh1 @ W1[:, :64]   [N,64]
//   dW3 += dl^T @ h2     dW2 += dh2^T @ h1     dW1 += dh1^T @ [x_pt | x_img]
#include <cublas_v2.h>
#include <cuda_bf16.h>
#include <dlfcn.h>

#include <mutex>

#include "common.cuh"

namespace tdssf {
namespace bwd {

constexpr int kCp = 64, kCi = 384, kK1 = 448, kN1 = 256, kN2 = 64, kC = 16;

// ---- cuBLAS, resolved at first use ----------------------------------------------------------------------------
struct Cublas {
  void* lib = nullptr;
  cublasStatus_t (*create)(cublasHandle_t*) = nullptr;
  cublasStatus_t (*set_stream)(cublasHandle_t, cudaStream_t) = nullptr;
  cublasStatus_t (*gemm_ex)(cublasHandle_t, cublasOperation_t, cublasOperation_t, int, int, int, const void*, const void*,
                            cudaDataType, int, const void*, cudaDataType, int, const void*, void*, cudaDataType, int,
                            cublasComputeType_t, cublasGemmAlgo_t) = nullptr;
  cublasHandle_t handle[64] = {};
  bool ok = false;
};

static Cublas& cublas() {
  static Cublas c;
  static std::once_flag once;
  std::call_once(once, [] {
    for (const char* name : {"libcublas.so.12", "libcublas.so", "/usr/local/cuda/lib64/libcublas.so.12"}) {
      c.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (c.lib) break;
    }
    if (!c.lib) return;
    c.create = reinterpret_cast<decltype(c.create)>(dlsym(c.lib, "cublasCreate_v2"));
    c.set_stream = reinterpret_cast<decltype(c.set_stream)>(dlsym(c.lib, "cublasSetStream_v2"));
    c.gemm_ex = reinterpret_cast<decltype(c.gemm_ex)>(dlsym(c.lib, "cublasGemmEx"));
    c.ok = c.create && c.set_stream && c.gemm_ex;
  });
  return c;
}

static cublasHandle_t handle_for_current_device() {
  static std::mutex mu;
  Cublas& c = cublas();
  if (!c.ok) return nullptr;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (!c.handle[dev] && c.create(&c.handle[dev]) != CUBLAS_STATUS_SUCCESS) c.handle[dev] = nullptr;
  return c.handle[dev];
}

// row-major C[m,n] (ldc) = alpha * op(A) op(B) + beta * C, A and B bf16 row-major, C fp32 (or bf16: c_bf16), fp32 accumulate
static cublasStatus_t gemm_rm(cublasHandle_t h, bool ta, bool tb, int m, int n, int k, float alpha, const void* A, int lda,
                              const void* B, int ldb, float beta, void* C, int ldc, bool c_bf16 = false) {
  // row-major X is column-major X^T: C^T = op(B)^T op(A)^T
  return cublas().gemm_ex(h, tb ? CUBLAS_OP_T : CUBLAS_OP_N, ta ? CUBLAS_OP_T : CUBLAS_OP_N, n, m, k, &alpha, B, CUDA_R_16BF,
                          ldb, A, CUDA_R_16BF, lda, &beta, C, c_bf16 ? CUDA_R_16BF : CUDA_R_32F, ldc, CUBLAS_COMPUTE_32F,
                          CUBLAS_GEMM_DEFAULT);
}

// ---- element-wise kernels -----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// dst[i] = bf16(src[i]), 8 elements per thread (every count is a multiple of 8);
// the three weight matrices in one launch (the training step is bound by the host's launches, not by this kernel)
__global__ void cvt3_bf16_kernel(const float* __restrict__ s0, uint4* __restrict__ d0, int64_t n0, const float* __restrict__ s1,
                                 uint4* __restrict__ d1, int64_t n1, const float* __restrict__ s2, uint4* __restrict__ d2, int64_t n2) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n0 + n1 + n2; i += (int64_t)gridDim.x * blockDim.x) {
    const float* src = i < n0 ? s0 : i < n0 + n1 ? s1 : s2;
    uint4* dst = i < n0 ? d0 : i < n0 + n1 ? d1 : d2;
    const int64_t k = i < n0 ? i : i < n0 + n1 ? i - n0 : i - n0 - n1;
    const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * k), b = __ldg(reinterpret_cast<const float4*>(src) + 2 * k + 1);
    dst[k] = make_uint4(pack2(a.x, a.y), pack2(a.z, a.w), pack2(b.x, b.y), pack2(b.z, b.w));
  }
}

// X[r, 0:64] = bf16(x_pt[r]), X[r, 64:448] = bf16(x_img[r]); one thread per 8 columns
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

// out = bf16(G * (h > 0)) (h == nullptr: no mask), db[c] += column sums of the masked values.  G is fp32 (the loss
// gradient) or bf16 (a data-gradient GEMM that wrote bf16 straight into `out`: in place, half the traffic of an fp32
// round trip, and the same bits - the mask is 0/1, so rounding before or after it is the same rounding).
// Block = 256 threads = 32 rows x 8 column groups per pass over a [N, C] matrix, C in {16, 64, 256}: thread (ty, tx)
// walks rows ty, ty+32, ... of its grid-stride slice and columns 8*tx + 64*j.
template <int C, typename GT>
__global__ void __launch_bounds__(256)
mask_cvt_colsum_kernel(const GT* G, const uint16_t* __restrict__ h, uint16_t* out, float* __restrict__ db, int64_t N) {
  constexpr int kGroups = C / 8;                      // 8-column groups per row
  constexpr int kTx = kGroups < 8 ? kGroups : 8;      // column groups handled side by side
  constexpr int kTy = 256 / kTx;
  constexpr int kJ = kGroups / kTx;
  const int tx = threadIdx.x % kTx, ty = threadIdx.x / kTx;
  float acc[kJ][8];
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[j][e] = 0.f;
  for (int64_t r = (int64_t)blockIdx.x * kTy + ty; r < N; r += (int64_t)gridDim.x * kTy) {
#pragma unroll
    for (int j = 0; j < kJ; ++j) {
      const int c = (tx + j * kTx) * 8;
      float v[8];
      if (sizeof(GT) == 4) {
        const float4* g4 = reinterpret_cast<const float4*>(G + r * C + c);
        const float4 a = g4[0], b = g4[1];
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
      } else {
        const uint4 q = *reinterpret_cast<const uint4*>(G + r * C + c);
        const uint32_t g[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float((e & 1) ? (g[e >> 1] & 0xffff0000u) : (g[e >> 1] << 16));
      }
      if (h) {
        const uint4 m = __ldg(reinterpret_cast<const uint4*>(h + r * C + c));
        const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const uint32_t bits = (w[e >> 1] >> ((e & 1) * 16)) & 0xffffu;   // bf16 > 0  <=>  sign clear and not zero
          if (bits == 0u || (bits & 0x8000u)) v[e] = 0.f;
        }
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[j][e] += v[e];
      *reinterpret_cast<uint4*>(out + r * C + c) = make_uint4(pack2(v[0], v[1]), pack2(v[2], v[3]), pack2(v[4], v[5]), pack2(v[6], v[7]));
    }
  }
  // lanes l and l + kTx, l + 2 kTx, ... of a warp hold the same columns: butterfly over them, then one shared-memory
  // atomic per warp and column, one global atomic per block and column
  __shared__ float s_sum[C];
  for (int c = threadIdx.x; c < C; c += 256) s_sum[c] = 0.f;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float v = acc[j][e];
#pragma unroll
      for (int o = 16; o >= kTx; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if ((threadIdx.x & 31) < kTx) atomicAdd(&s_sum[(tx + j * kTx) * 8 + e], v);
    }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 256) atomicAdd(db + c, s_sum[c]);
}

struct Work {
  uint16_t *X, *dl, *dh2, *dh1, *W1, *W2, *W3;
};
static size_t al(size_t b) { return (b + 255) & ~(size_t)255; }
static size_t carve(int64_t N, char* base, Work* w) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += al(bytes); return p; };
  uint16_t* X = (uint16_t*)take((size_t)N * kK1 * 2);
  uint16_t* dl = (uint16_t*)take((size_t)N * kC * 2);
  uint16_t* dh2 = (uint16_t*)take((size_t)N * kN2 * 2);
  uint16_t* dh1 = (uint16_t*)take((size_t)N * kN1 * 2);
  uint16_t* W1 = (uint16_t*)take((size_t)kN1 * kK1 * 2);
  uint16_t* W2 = (uint16_t*)take((size_t)kN2 * kN1 * 2);
  uint16_t* W3 = (uint16_t*)take((size_t)kC * kN2 * 2);
  if (w) *w = Work{X, dl, dh2, dh1, W1, W2, W3};
  return off;
}

}  // namespace bwd
}  // namespace tdssf

extern "C" size_t tdssf_mlp_bwd_tc_workspace_bytes(int64_t N) {
  return tdssf::bwd::carve(N < 0 ? 0 : N, nullptr, nullptr) + 256;
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
  TDSSF_REQUIRE(N >= 0 && N < (1ll << 31), "tdssf_mlp_bwd_tc: N out of range");
  if (N == 0) return TDSSF_OK;
  TDSSF_REQUIRE((((uintptr_t)x_pt | (uintptr_t)x_img | (uintptr_t)x_bf16 | (uintptr_t)h1_bf16 | (uintptr_t)h2_bf16 | (uintptr_t)dlogits |
                  (uintptr_t)dx_pt) & 15) == 0, "tdssf_mlp_bwd_tc: pointers must be 16-byte aligned");
  TDSSF_REQUIRE(workspace && ((uintptr_t)workspace & 255) == 0 && workspace_bytes >= tdssf_mlp_bwd_tc_workspace_bytes(N),
                "tdssf_mlp_bwd_tc: 256-byte aligned workspace of %zu bytes needed, got %zu",
                tdssf_mlp_bwd_tc_workspace_bytes(N), workspace_bytes);
  if (int rc = tdssf_check_device()) return rc;
  cublasHandle_t h = handle_for_current_device();
  if (!h) {
    set_error("tdssf_mlp_bwd_tc: cuBLAS (libcublas.so.12) could not be loaded");
    return TDSSF_ENOSUP;
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (cublas().set_stream(h, s) != CUBLAS_STATUS_SUCCESS) {
    set_error("tdssf_mlp_bwd_tc: cublasSetStream failed");
    return TDSSF_ECUDA;
  }
  Work w;
  carve(N, reinterpret_cast<char*>(workspace), &w);
  const unsigned blocks = (unsigned)num_sms() * 8;
  // operands in bf16
  cvt3_bf16_kernel<<<blocks, 256, 0, s>>>(w_host->W1, reinterpret_cast<uint4*>(w.W1), (int64_t)kN1 * kK1 / 8,
                                          w_host->W2, reinterpret_cast<uint4*>(w.W2), (int64_t)kN2 * kN1 / 8,
                                          w_host->W3, reinterpret_cast<uint4*>(w.W3), (int64_t)kC * kN2 / 8);
  const void* X = x_bf16;
  if (!X) {
    concat_bf16_kernel<<<blocks, 256, 0, s>>>(x_pt, x_img, reinterpret_cast<uint4*>(w.X), N);
    X = w.X;
  }
  mask_cvt_colsum_kernel<kC, float><<<blocks, 256, 0, s>>>(dlogits, nullptr, w.dl, g_host->db3, N);
  TDSSF_LAUNCH_CHECK();
  const int n = (int)N;
  auto ok = [&](cublasStatus_t st, const char* what) {
    if (st == CUBLAS_STATUS_SUCCESS) return true;
    set_error("tdssf_mlp_bwd_tc: cublasGemmEx(%s) failed with status %d", what, (int)st);
    return false;
  };
  // data gradients
  // (the two hidden-layer data gradients leave cuBLAS as bf16, directly in the buffers the mask kernels finish in place)
  if (!ok(gemm_rm(h, false, false, n, kN2, kC, 1.f, w.dl, kC, w.W3, kN2, 0.f, w.dh2, kN2, true), "dH2")) return TDSSF_ECUDA;
  mask_cvt_colsum_kernel<kN2, uint16_t><<<blocks, 256, 0, s>>>(w.dh2, reinterpret_cast<const uint16_t*>(h2_bf16), w.dh2, g_host->db2, N);
  if (!ok(gemm_rm(h, false, false, n, kN1, kN2, 1.f, w.dh2, kN2, w.W2, kN1, 0.f, w.dh1, kN1, true), "dH1")) return TDSSF_ECUDA;
  mask_cvt_colsum_kernel<kN1, uint16_t><<<blocks, 256, 0, s>>>(w.dh1, reinterpret_cast<const uint16_t*>(h1_bf16), w.dh1, g_host->db1, N);
  TDSSF_LAUNCH_CHECK();
  if (dx_pt && !ok(gemm_rm(h, false, false, n, kCp, kN1, 1.f, w.dh1, kN1, w.W1, kK1, 0.f, dx_pt, kCp), "dx_pt")) return TDSSF_ECUDA;
  // weight gradients, accumulated into the caller's fp32 buffers
  if (!ok(gemm_rm(h, true, false, kC, kN2, n, 1.f, w.dl, kC, h2_bf16, kN2, 1.f, g_host->dW3, kN2), "dW3")) return TDSSF_ECUDA;
  if (!ok(gemm_rm(h, true, false, kN2, kN1, n, 1.f, w.dh2, kN2, h1_bf16, kN1, 1.f, g_host->dW2, kN1), "dW2")) return TDSSF_ECUDA;
  if (!ok(gemm_rm(h, true, false, kN1, kK1, n, 1.f, w.dh1, kN1, X, kK1, 1.f, g_host->dW1, kK1), "dW1")) return TDSSF_ECUDA;
  return TDSSF_OK;
}
