// a11: confusion matrix (utils/evaluation.py:35-42): cm[C*(label-1) + (pred-1)] += 1, int64.
// Shared-memory privatised histogram per CTA, one 64-bit atomic per non-empty bin per CTA.
// HBM-bound on the 16 B/pair input; the matrix itself is 2 KB.
#include <algorithm>
#include "common.cuh"

namespace tdssf {

constexpr int kCmThreads = 256;
constexpr int kCmMaxBins = 64 * 64;

__global__ void __launch_bounds__(kCmThreads)
confusion_kernel(const int64_t* __restrict__ preds, const int64_t* __restrict__ labels, int64_t n, int C,
                 unsigned long long* __restrict__ cm) {
  extern __shared__ unsigned int s_bins[];
  const int bins = C * C;
  for (int i = threadIdx.x; i < bins; i += blockDim.x) s_bins[i] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = __ldg(preds + i) - 1, l = __ldg(labels + i) - 1;
    // torch.bincount would raise on a negative index and grow on a large one; out-of-range pairs
    // are dropped here and rejected on the host side of the wrapper.
    if (p >= 0 && p < C && l >= 0 && l < C) atomicAdd(&s_bins[(int)l * C + (int)p], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < bins; i += blockDim.x)
    if (s_bins[i]) atomicAdd(cm + i, (unsigned long long)s_bins[i]);
}

}  // namespace tdssf

extern "C" int tdssf_confusion(const int64_t* preds, const int64_t* labels, int64_t n, int C, int64_t* cm,
                               void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(cm, "tdssf_confusion: NULL cm");
  TDSSF_REQUIRE(C >= 1 && C * C <= kCmMaxBins, "tdssf_confusion: C=%d not in 1..64", C);
  TDSSF_REQUIRE(n >= 0, "tdssf_confusion: negative n");
  if (n == 0) return TDSSF_OK;
  TDSSF_REQUIRE(preds && labels, "tdssf_confusion: NULL input");
  // each CTA's 32-bit private bins cannot overflow: a CTA sees at most n / gridDim + blockDim pairs
  const int64_t blocks = std::min<int64_t>(ceil_div(n, kCmThreads * 8), (int64_t)num_sms() * 8);
  confusion_kernel<<<(unsigned)std::max<int64_t>(blocks, 1), kCmThreads, C * C * sizeof(unsigned int),
                     (cudaStream_t)stream>>>(preds, labels, n, C, reinterpret_cast<unsigned long long*>(cm));
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}
