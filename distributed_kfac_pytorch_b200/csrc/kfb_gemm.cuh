// Grouped "TN" GEMM shared by the SIMT and tcgen05 back ends:
//   D[m][n] = sum_k A[m][k] * B[n][k]        (A, B, D row-major fp32 with pitches)
//   epilogue: optional elementwise multiplier Mul[m][n]; optional dot-product with Dot[m][n]
//   accumulated into *dot_acc (used for the KL-clip inner product).
#pragma once
#include "kfb_common.cuh"

namespace kfb {

struct GemmProb {
  const float* A;
  const float* B;
  float* D;
  const float* Mul;
  const float* Dot;
  int M, N, K;
  int pad_;
  int64_t lda, ldb, ldd, ldmul, lddot;
};

constexpr int kGemmBatch = 128;
typedef Batch<GemmProb, kGemmBatch> GemmBatch;

// Run `n` problems (one launch per kGemmBatch problems). SIMT fp32.
KFB_HIDDEN int gemm_grouped_simt(const GemmProb* probs, int n, float* dot_acc, cudaStream_t stream);
// tcgen05 TF32; problems must satisfy gemm_tc_eligible().
KFB_HIDDEN int gemm_grouped_tc(Context* ctx, const GemmProb* probs, int n, float* dot_acc, cudaStream_t stream, bool x3 = false);
KFB_HIDDEN bool gemm_tc_eligible(const GemmProb& p);

}  // namespace kfb
