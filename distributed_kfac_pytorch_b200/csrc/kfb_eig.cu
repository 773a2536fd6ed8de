// Eigendecomposition (cuSOLVER syevd: the one library call of the path) and the small
// kernels around it: eigenvalue clip, transpose, outer-product reciprocal.
#include <cusolverDn.h>

#include <algorithm>

#include "kfb_common.cuh"

namespace kfb {
namespace {

__global__ void clip_kernel(float* w, int d) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < d) w[i] = fmaxf(w[i], 0.f);
}

// out[c][r] = in[r][c]; 32x32 tiles through shared memory, both sides coalesced
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ in, int64_t ld_in,
                                                       int rows, int cols, float* __restrict__ out,
                                                       int64_t ld_out) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    int r = r0 + ty + k * 8, c = c0 + tx;
    tile[ty + k * 8][tx] = (r < rows && c < cols) ? in[(int64_t)r * ld_in + c] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    int c = c0 + ty + k * 8, r = r0 + tx;
    if (c < cols && r < rows) out[(int64_t)c * ld_out + r] = tile[tx][ty + k * 8];
  }
}

__global__ void __launch_bounds__(256) outer_recip_kernel(const float* __restrict__ dG,
                                                         const float* __restrict__ dA, int dg, int da,
                                                         float damping, float* __restrict__ out,
                                                         int64_t ld) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= da) return;
  const float a = dA[j];
  for (int i = blockIdx.y; i < dg; i += gridDim.y) out[(int64_t)i * ld + j] = 1.f / (dG[i] * a + damping);
}

int launch_transpose(const float* in, int64_t ld_in, int rows, int cols, float* out, int64_t ld_out,
                     cudaStream_t stream) {
  dim3 grid(ceil_div(cols, 32), ceil_div(rows, 32));
  transpose_kernel<<<grid, 256, 0, stream>>>(in, ld_in, rows, cols, out, ld_out);
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

}  // namespace
}  // namespace kfb

extern "C" {

size_t kfb_syevd_workspace_bytes(int32_t d) {
  kfb::Context* ctx = kfb::current_context();
  if (!ctx || d <= 0) return 0;
  int lwork = 0;
  // the query needs no valid pointers
  cusolverStatus_t st = cusolverDnSsyevd_bufferSize((cusolverDnHandle_t)ctx->cusolver, CUSOLVER_EIG_MODE_VECTOR,
                                                    CUBLAS_FILL_MODE_UPPER, d, nullptr, (d + 3) / 4 * 4,
                                                    nullptr, &lwork);
  if (st != CUSOLVER_STATUS_SUCCESS) return 0;
  return (size_t)lwork * sizeof(float) + 256;
}

int kfb_syevd(const float* factor, int64_t ld_factor, int32_t d, float* E, int64_t ld_e, float* Q,
              int64_t ld_q, float* w, void* workspace, size_t workspace_bytes, int32_t* dev_info,
              void* stream_) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(factor && E && Q && w && dev_info && d > 0, "null argument");
  KFB_REQUIRE(ld_factor >= d && ld_e >= d && ld_q >= d, "pitch smaller than d");
  cudaStream_t stream = (cudaStream_t)stream_;
  cusolverDnHandle_t h = (cusolverDnHandle_t)ctx->cusolver;
  if (cusolverDnSetStream(h, stream) != CUSOLVER_STATUS_SUCCESS) KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnSetStream failed");
  int lwork = 0;
  if (cusolverDnSsyevd_bufferSize(h, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, d, E, (int)ld_e, w, &lwork) !=
      CUSOLVER_STATUS_SUCCESS)
    KFB_FAIL(KFB_ERR_CUSOLVER, "syevd bufferSize failed for d=%d", d);
  if (!workspace || workspace_bytes < (size_t)lwork * sizeof(float))
    KFB_FAIL(KFB_ERR_WORKSPACE, "syevd workspace: need %zu bytes, got %zu", (size_t)lwork * sizeof(float), workspace_bytes);
  // syevd overwrites its input with the eigenvectors.  The factor is symmetric, so the
  // row-major copy IS the column-major matrix; on exit column j (column-major) = eigenvector j,
  // i.e. row j of the row-major view: E = Q^T comes for free, Q needs one transpose.
  KFB_CUDA(cudaMemcpy2DAsync(E, ld_e * sizeof(float), factor, ld_factor * sizeof(float), (size_t)d * sizeof(float), d,
                             cudaMemcpyDeviceToDevice, stream));
  cusolverStatus_t st = cusolverDnSsyevd(h, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, d, E, (int)ld_e, w,
                                         (float*)workspace, lwork, dev_info);
  if (st != CUSOLVER_STATUS_SUCCESS) KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnSsyevd failed (%d) for d=%d", (int)st, d);
  kfb::clip_kernel<<<kfb::ceil_div(d, 256), 256, 0, stream>>>(w, d);
  KFB_CHECK_LAUNCH();
  return kfb::launch_transpose(E, ld_e, d, d, Q, ld_q, stream);
}

int kfb_transpose(const float* in, int64_t ld_in, int32_t rows, int32_t cols, float* out, int64_t ld_out,
                  void* stream) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(in && out && rows > 0 && cols > 0 && ld_in >= cols && ld_out >= rows, "bad transpose arguments");
  return kfb::launch_transpose(in, ld_in, rows, cols, out, ld_out, (cudaStream_t)stream);
}

int kfb_outer_recip(const float* dG, const float* dA, int32_t dg, int32_t da, float damping, float* out,
                    int64_t ld_out, void* stream) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(dG && dA && out && dg > 0 && da > 0 && ld_out >= da, "bad outer_recip arguments");
  dim3 grid(kfb::ceil_div(da, 256), std::min(dg, 256));
  kfb::outer_recip_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dG, dA, dg, da, damping, out, ld_out);
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

}  // extern "C"
