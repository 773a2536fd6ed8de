// Eigendecomposition (cuSOLVER syevd: the one library call of the path) and the small
// kernels around it: eigenvalue clip, transpose, outer-product reciprocal.
#include <cusolverDn.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

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

static int syevd_one(cusolverDnHandle_t h, const float* factor, int64_t ld_factor, int32_t d, float* E, int64_t ld_e,
                     float* Q, int64_t ld_q, float* w, void* workspace, size_t workspace_bytes, int32_t* dev_info,
                     cudaStream_t stream) {
  KFB_REQUIRE(factor && E && Q && w && dev_info && d > 0, "null argument");
  KFB_REQUIRE(ld_factor >= d && ld_e >= d && ld_q >= d, "pitch smaller than d");
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

int kfb_syevd(const float* factor, int64_t ld_factor, int32_t d, float* E, int64_t ld_e, float* Q,
              int64_t ld_q, float* w, void* workspace, size_t workspace_bytes, int32_t* dev_info,
              void* stream_) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  return syevd_one((cusolverDnHandle_t)ctx->cusolver, factor, ld_factor, d, E, ld_e, Q, ld_q, w, workspace,
                   workspace_bytes, dev_info, (cudaStream_t)stream_);
}

// ---- batched: worker lanes -------------------------------------------------------------------------
namespace {
constexpr int kMaxLanes = 8;
struct Lane {
  cudaStream_t stream = nullptr;
  cusolverDnHandle_t handle = nullptr;
  cudaEvent_t done = nullptr;
};
struct LanePool {
  Lane lanes[kMaxLanes];
  cudaEvent_t start = nullptr;
  bool ready = false;
};
LanePool g_pools[64];
std::mutex g_pool_mu;

int get_pool(kfb::Context* ctx, LanePool** out) {
  std::lock_guard<std::mutex> lock(g_pool_mu);
  LanePool& p = g_pools[ctx->device];
  if (!p.ready) {
    KFB_CUDA(cudaEventCreateWithFlags(&p.start, cudaEventDisableTiming));
    for (int i = 0; i < kMaxLanes; ++i) {
      KFB_CUDA(cudaStreamCreateWithFlags(&p.lanes[i].stream, cudaStreamNonBlocking));
      KFB_CUDA(cudaEventCreateWithFlags(&p.lanes[i].done, cudaEventDisableTiming));
      if (cusolverDnCreate(&p.lanes[i].handle) != CUSOLVER_STATUS_SUCCESS)
        KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnCreate failed for worker lane %d", i);
    }
    p.ready = true;
  }
  *out = &p;
  return KFB_OK;
}

// longest-processing-time placement of the jobs on the lanes; returns per-lane job lists
void place_jobs(const kfb_syevd_job* jobs, int n, std::vector<std::vector<int>>* lanes) {
  int n_lanes = std::min(kMaxLanes, std::max(1, n));
  lanes->assign(n_lanes, {});
  std::vector<int> order(n);
  for (int i = 0; i < n; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return jobs[a].d > jobs[b].d; });
  std::vector<double> load(n_lanes, 0.0);
  for (int i : order) {
    int best = 0;
    for (int l = 1; l < n_lanes; ++l)
      if (load[l] < load[best]) best = l;
    double d = jobs[i].d;
    load[best] += d * d * d + 2e6;      // + a constant: tiny solves are pure launch latency
    (*lanes)[best].push_back(i);
  }
}

size_t lwork_bytes(cusolverDnHandle_t h, int d) {
  int lwork = 0;
  if (cusolverDnSsyevd_bufferSize(h, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, d, nullptr, (d + 3) / 4 * 4, nullptr,
                                  &lwork) != CUSOLVER_STATUS_SUCCESS)
    return 0;
  return ((size_t)lwork * sizeof(float) + 255) / 256 * 256 + 256;
}
}  // namespace

size_t kfb_syevd_batch_workspace_bytes(const kfb_syevd_job* jobs, int32_t n_jobs) {
  kfb::Context* ctx = kfb::current_context();
  if (!ctx || !jobs || n_jobs <= 0) return 0;
  std::vector<std::vector<int>> lanes;
  place_jobs(jobs, n_jobs, &lanes);
  size_t total = 0;
  for (auto& l : lanes) {
    size_t mx = 0;
    for (int i : l) mx = std::max(mx, lwork_bytes((cusolverDnHandle_t)ctx->cusolver, jobs[i].d));
    total += mx;
  }
  return total;
}

int kfb_syevd_batch(const kfb_syevd_job* jobs, int32_t n_jobs, void* workspace, size_t workspace_bytes,
                    void* stream_) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(n_jobs >= 0 && (jobs || n_jobs == 0), "bad job array");
  if (n_jobs == 0) return KFB_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  LanePool* pool;
  if ((rc = get_pool(ctx, &pool))) return rc;
  std::vector<std::vector<int>> lanes;
  place_jobs(jobs, n_jobs, &lanes);
  const int n_lanes = (int)lanes.size();
  std::vector<size_t> off(n_lanes + 1, 0);
  for (int l = 0; l < n_lanes; ++l) {
    size_t mx = 0;
    for (int i : lanes[l]) mx = std::max(mx, lwork_bytes((cusolverDnHandle_t)ctx->cusolver, jobs[i].d));
    off[l + 1] = off[l] + mx;
  }
  if (!workspace || workspace_bytes < off[n_lanes])
    KFB_FAIL(KFB_ERR_WORKSPACE, "syevd batch workspace: need %zu bytes, got %zu", off[n_lanes], workspace_bytes);
  KFB_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  // fork: every lane starts after what is already queued on the caller's stream
  KFB_CUDA(cudaEventRecord(pool->start, stream));
  std::vector<int> status(n_lanes, KFB_OK);
  std::vector<std::string> errors(n_lanes);
  std::vector<std::thread> threads;
  const int device = ctx->device;
  for (int l = 0; l < n_lanes; ++l) {
    threads.emplace_back([&, l]() {
      // cuSOLVER's syevd synchronises with the host internally, so each lane needs its own thread
      // for the lanes to overlap at all
      cudaSetDevice(device);
      kfb_init(device);
      Lane& lane = pool->lanes[l];
      if (cudaStreamWaitEvent(lane.stream, pool->start, 0) != cudaSuccess) { status[l] = KFB_ERR_CUDA; return; }
      for (int i : lanes[l]) {
        const kfb_syevd_job& j = jobs[i];
        int r = syevd_one(lane.handle, j.factor, j.ld_factor, j.d, j.E, j.ld_e, j.Q, j.ld_q, j.w,
                          (char*)workspace + off[l], off[l + 1] - off[l], j.dev_info, lane.stream);
        if (r) { status[l] = r; errors[l] = kfb::get_error(); break; }
      }
      if (cudaEventRecord(lane.done, lane.stream) != cudaSuccess) status[l] = KFB_ERR_CUDA;
    });
  }
  for (auto& t : threads) t.join();
  // join: the caller's stream continues after every lane
  for (int l = 0; l < n_lanes; ++l) KFB_CUDA(cudaStreamWaitEvent(stream, pool->lanes[l].done, 0));
  for (int l = 0; l < n_lanes; ++l)
    if (status[l]) KFB_FAIL(status[l], "syevd batch lane %d failed: %s", l, errors[l].c_str());
  return KFB_OK;
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
