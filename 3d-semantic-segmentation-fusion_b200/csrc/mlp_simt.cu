// a6, fp32 CUDA-core path of the fusion MLP (the 1e-5 parity mode; the tensor-core path is mlp_tc.cu).
// Replaces utils/fusion_model.py:144-150,203-204 and its autograd backward.
//   forward : H1 = relu([Xpt|Ximg] W1^T + b1), H2 = relu(H1 W2^T + b2), logits = H2 W3^T + b3
//   backward: dW*, db* accumulated (+=), dXpt written; no gradient for the image feature
// All GEMMs use one 64x64x16 shared-memory tile with a 4x4 register micro-tile per thread.
#include <algorithm>
#include "common.cuh"

namespace tdssf {

constexpr int kBM = 64, kBN = 64, kBK = 16, kSimtThreads = 256;

// A operand = rows of [A0 | A1] (A1 may be NULL / K1 = 0), row-major, K0 and K1 multiples of kBK.
struct ARows {
  const float* a0; int k0;
  const float* a1; int k1;
};

enum { kLayoutNK = 0 /* B[n][k] row-major: C = A B^T */, kLayoutKN = 1 /* B[k][n] row-major: C = A B */ };

// C[m, n] = epi( sum_k A[m,k] * Bop[k,n] )      m < M, n < N
//   EPI 0: + bias[n]            EPI 1: relu(+ bias[n])        EPI 2: * (gate[m,n] > 0)
template <int BLAYOUT, int EPI>
__global__ void __launch_bounds__(kSimtThreads)
gemm_rows_kernel(ARows A, int64_t M, const float* __restrict__ Bm, int ldb, int N, int Ktot,
                 const float* __restrict__ bias, const float* __restrict__ gate,
                 float* __restrict__ C, int ldc, int ny) {
  __shared__ float As[kBK][kBM + 4];
  __shared__ float Bs[kBK][kBN + 4];
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  // 1-D grid, column block fastest: the `ny` CTAs that share a block of rows run together and read it from DRAM once
  // (with the column block in blockIdx.y every column block re-read its rows: 1.99 GB for 0.50 GB of operands in ncu)
  const int64_t m0 = (int64_t)(blockIdx.x / (unsigned)ny) * kBM;
  const int n0 = (int)(blockIdx.x % (unsigned)ny) * kBN;
  float acc[4][4] = {};

  for (int kk = 0; kk < Ktot; kk += kBK) {
    {  // A tile: 64 rows x 16 k, one float4 along k per thread
      const int r = t >> 2, kq = (t & 3) * 4;
      const int64_t m = m0 + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < M) {
        const float* src = (kk < A.k0) ? (A.a0 + m * A.k0 + kk + kq) : (A.a1 + m * A.k1 + (kk - A.k0) + kq);
        v = __ldg(reinterpret_cast<const float4*>(src));
      }
      As[kq + 0][r] = v.x; As[kq + 1][r] = v.y; As[kq + 2][r] = v.z; As[kq + 3][r] = v.w;
    }
    if (BLAYOUT == kLayoutNK) {  // B[n][k]
      const int r = t >> 2, kq = (t & 3) * 4;
      const int n = n0 + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n < N) v = __ldg(reinterpret_cast<const float4*>(Bm + (int64_t)n * ldb + kk + kq));
      Bs[kq + 0][r] = v.x; Bs[kq + 1][r] = v.y; Bs[kq + 2][r] = v.z; Bs[kq + 3][r] = v.w;
    } else {  // B[k][n]
      const int k = t >> 4, nq = (t & 15) * 4;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int n = n0 + nq + i;
        Bs[k][nq + i] = (n < N) ? __ldg(Bm + (int64_t)(kk + k) * ldb + n) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kBK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[k][ty * 4 + i]; b[i] = Bs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j];
      if (EPI == 0) v += bias[n];
      if (EPI == 1) v = fmaxf(v + bias[n], 0.f);
      if (EPI == 2) v = (gate[m * ldc + n] > 0.f) ? v : 0.f;
      C[m * ldc + n] = v;
    }
  }
}

// Weight gradient: dW[o, i] += sum_m dY[m, o] * X[m, i] over a slab of rows; db[o] += sum_m dY[m,o].
// grid = (ceil(O/64), ceil(I/64), slabs); split-K over rows with fp32 atomics.
__global__ void __launch_bounds__(kSimtThreads)
wgrad_kernel(const float* __restrict__ dY, int O, ARows X, int I, int64_t M, int64_t rows_per_slab,
             float* __restrict__ dW, float* __restrict__ db) {
  __shared__ float Ys[kBK][kBM + 4];  // [m][o]
  __shared__ float Xs[kBK][kBN + 4];  // [m][i]
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int o0 = blockIdx.x * kBM, i0 = blockIdx.y * kBN;
  const int64_t mbeg = (int64_t)blockIdx.z * rows_per_slab;
  const int64_t mend = min(M, mbeg + rows_per_slab);
  float acc[4][4] = {};
  float bacc[4] = {};
  for (int64_t mm = mbeg; mm < mend; mm += kBK) {
    {
      const int r = t >> 4, q = (t & 15) * 4;
      const int64_t m = mm + r;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int o = o0 + q + i;
        Ys[r][q + i] = (m < mend && o < O) ? __ldg(dY + m * O + o) : 0.f;
      }
      const int ci = i0 + q;  // 4 consecutive input columns, never straddling the concat boundary
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < mend && ci < I) {
        const float* src = (ci < X.k0) ? (X.a0 + m * X.k0 + ci) : (X.a1 + m * X.k1 + (ci - X.k0));
        v = __ldg(reinterpret_cast<const float4*>(src));
      }
      Xs[r][q + 0] = v.x; Xs[r][q + 1] = v.y; Xs[r][q + 2] = v.z; Xs[r][q + 3] = v.w;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kBK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = Ys[k][ty * 4 + i]; b[i] = Xs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (tx == 0) bacc[i] += a[i];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = o0 + ty * 4 + i;
    if (o >= O) continue;
    if (tx == 0 && blockIdx.y == 0 && db) atomicAdd(db + o, bacc[i]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ii = i0 + tx * 4 + j;
      if (ii < I) atomicAdd(dW + (int64_t)o * I + ii, acc[i][j]);
    }
  }
}

static int check_weights(const tdssf_mlp_weights* w, const char* who) {
  TDSSF_REQUIRE(w && w->W1 && w->b1 && w->W2 && w->b2 && w->W3 && w->b3, "%s: NULL weight pointer", who);
  return TDSSF_OK;
}

}  // namespace tdssf

extern "C" int tdssf_mlp_fwd_f32(const float* x_pt, const float* x_img, int64_t N,
                                 const tdssf_mlp_weights* w, float* h1, float* h2, float* logits,
                                 void* stream) {
  using namespace tdssf;
  if (int rc = check_weights(w, "tdssf_mlp_fwd_f32")) return rc;
  TDSSF_REQUIRE(x_pt && x_img && logits && h1 && h2, "tdssf_mlp_fwd_f32: NULL pointer (h1/h2 scratch is required)");
  TDSSF_REQUIRE(N >= 0, "tdssf_mlp_fwd_f32: negative N");
  TDSSF_REQUIRE((((uintptr_t)x_pt | (uintptr_t)x_img | (uintptr_t)h1 | (uintptr_t)h2 | (uintptr_t)w->W1 |
                  (uintptr_t)w->W2 | (uintptr_t)w->W3) & 15) == 0,
                "tdssf_mlp_fwd_f32: operands must be 16-byte aligned");
  if (N == 0) return TDSSF_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const unsigned gm = (unsigned)ceil_div(N, kBM);
  ARows a1{x_pt, 64, x_img, 384};
  gemm_rows_kernel<kLayoutNK, 1><<<dim3(gm * 4), kSimtThreads, 0, s>>>(a1, N, w->W1, 448, 256, 448, w->b1, nullptr, h1, 256, 4);
  TDSSF_LAUNCH_CHECK();
  ARows a2{h1, 256, nullptr, 0};
  gemm_rows_kernel<kLayoutNK, 1><<<dim3(gm * 1), kSimtThreads, 0, s>>>(a2, N, w->W2, 256, 64, 256, w->b2, nullptr, h2, 64, 1);
  TDSSF_LAUNCH_CHECK();
  ARows a3{h2, 64, nullptr, 0};
  gemm_rows_kernel<kLayoutNK, 0><<<dim3(gm * 1), kSimtThreads, 0, s>>>(a3, N, w->W3, 64, 16, 64, w->b3, nullptr, logits, 16, 1);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

extern "C" size_t tdssf_mlp_bwd_workspace_bytes(int64_t N) {
  if (N < 0) N = 0;
  return (size_t)N * (256 + 64) * sizeof(float) + 256;  // dH1, dH2
}

extern "C" int tdssf_mlp_bwd_f32(const float* x_pt, const float* x_img, const float* h1, const float* h2,
                                 const float* dlogits, int64_t N, const tdssf_mlp_weights* w,
                                 const tdssf_mlp_grads* g, float* dx_pt, void* workspace,
                                 size_t workspace_bytes, void* stream) {
  using namespace tdssf;
  if (int rc = check_weights(w, "tdssf_mlp_bwd_f32")) return rc;
  TDSSF_REQUIRE(g && g->dW1 && g->db1 && g->dW2 && g->db2 && g->dW3 && g->db3, "tdssf_mlp_bwd_f32: NULL gradient pointer");
  TDSSF_REQUIRE(x_pt && x_img && h1 && h2 && dlogits, "tdssf_mlp_bwd_f32: NULL input");
  TDSSF_REQUIRE(N >= 0, "tdssf_mlp_bwd_f32: negative N");
  TDSSF_REQUIRE(workspace_bytes >= tdssf_mlp_bwd_workspace_bytes(N) && workspace, "tdssf_mlp_bwd_f32: workspace too small");
  if (N == 0) return TDSSF_OK;
  cudaStream_t s = (cudaStream_t)stream;
  float* dH1 = reinterpret_cast<float*>(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  float* dH2 = dH1 + N * 256;
  const unsigned gm = (unsigned)ceil_div(N, kBM);
  // slabs of rows for the split-K weight gradients: enough CTAs to fill the machine
  const int64_t slabs = std::max<int64_t>(1, std::min<int64_t>(ceil_div(N, 2048), 4 * num_sms()));
  const int64_t rows_per_slab = ceil_div(ceil_div(N, slabs), kBK) * kBK;
  const unsigned gz = (unsigned)ceil_div(N, rows_per_slab);

  // layer 3
  ARows h2r{h2, 64, nullptr, 0};
  wgrad_kernel<<<dim3(1, 1, gz), kSimtThreads, 0, s>>>(dlogits, 16, h2r, 64, N, rows_per_slab, g->dW3, g->db3);
  TDSSF_LAUNCH_CHECK();
  ARows dl{dlogits, 16, nullptr, 0};
  gemm_rows_kernel<kLayoutKN, 2><<<dim3(gm * 1), kSimtThreads, 0, s>>>(dl, N, w->W3, 64, 64, 16, nullptr, h2, dH2, 64, 1);
  TDSSF_LAUNCH_CHECK();
  // layer 2
  ARows h1r{h1, 256, nullptr, 0};
  wgrad_kernel<<<dim3(1, 4, gz), kSimtThreads, 0, s>>>(dH2, 64, h1r, 256, N, rows_per_slab, g->dW2, g->db2);
  TDSSF_LAUNCH_CHECK();
  ARows dh2{dH2, 64, nullptr, 0};
  gemm_rows_kernel<kLayoutKN, 2><<<dim3(gm * 4), kSimtThreads, 0, s>>>(dh2, N, w->W2, 256, 256, 64, nullptr, h1, dH1, 256, 4);
  TDSSF_LAUNCH_CHECK();
  // layer 1
  ARows xr{x_pt, 64, x_img, 384};
  wgrad_kernel<<<dim3(4, 7, gz), kSimtThreads, 0, s>>>(dH1, 256, xr, 448, N, rows_per_slab, g->dW1, g->db1);
  TDSSF_LAUNCH_CHECK();
  if (dx_pt) {
    // dXpt = dH1 @ W1[:, :64]   (no activation gate: reuse EPI 0 with a zero bias is avoided by EPI 3)
    ARows dh1{dH1, 256, nullptr, 0};
    gemm_rows_kernel<kLayoutKN, 3><<<dim3(gm * 1), kSimtThreads, 0, s>>>(dh1, N, w->W1, 448, 64, 256, nullptr, nullptr, dx_pt, 64, 1);
    TDSSF_LAUNCH_CHECK();
  }
  return TDSSF_OK;
}
