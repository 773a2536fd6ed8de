// Eigendecomposition (cuSOLVER syevd: the one library call of the path) and the small
// kernels around it: eigenvalue clip, transpose, outer-product reciprocal.
#include <cuda.h>
#include <cusolverDn.h>

#include <stdlib.h>

#include <algorithm>
#include <cmath>
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

// ---- many decompositions at once ---------------------------------------------------------------
// Two mechanisms, both cuSOLVER:
//  * same-size groups of small/medium factors go through cusolverDnXsyevBatched (one launch
//    sequence for the whole group: 22 64x64 solves take 0.4 ms instead of 20 ms);
//  * the remaining (large or one-of-a-kind) factors run syevd on worker lanes, one host thread +
//    stream + handle each, longest first.
namespace {
constexpr int kMaxLanes = 8;
struct Lane {
  cudaStream_t stream = nullptr;
  cusolverDnHandle_t handle = nullptr;
  cudaEvent_t done = nullptr;
};
struct LanePool {
  Lane lanes[kMaxLanes];
  Lane grouped;                 // stream/handle for the XsyevBatched groups
  cusolverDnParams_t params = nullptr;
  cudaEvent_t start = nullptr;
  int n_green = 0;              // lanes [0, n_green) run in green contexts
  bool ready = false;
};
LanePool g_pools[64];
std::mutex g_pool_mu;

// How many syevd lanes to run: KFB_SYEVD_LANES, else three - one host thread per lane next to this
// process's main thread, fewer when the node's cores are shared with many ranks (LOCAL_WORLD_SIZE,
// set by torchrun).  With green contexts the lanes also split the SMs, and the large factors need
// SMs: ResNet-50 on one B200 takes 646 ms on ordinary streams and 560 / 434 / 491 / 608 / 518 / 737
// ms with 2 / 3 / 4 / 5 / 6 / 8 green-context lanes (64 / 48 / 32 / 24 / 24 / 16 SMs each).
int lane_count() {
  static const int n = []() {
    if (const char* e = getenv("KFB_SYEVD_LANES")) return std::min(kMaxLanes, std::max(1, atoi(e)));
    int local_world = 1;
    if (const char* e = getenv("LOCAL_WORLD_SIZE")) local_world = std::max(1, atoi(e));
    const int cores = (int)std::max(1u, std::thread::hardware_concurrency());
    return std::min(3, std::max(2, cores / local_world - 1));
  }();
  return n;
}

// Streams for the syevd lanes inside GREEN CONTEXTS (CUDA >= 12.4): each lane gets its own context
// with a slice of the SMs.  syevd is bound by its host-side panel loop (thousands of microsecond
// kernels), and launches into ONE context serialise across threads - two lanes on ordinary streams
// reach 1.33x, more add nothing - while launches into different green contexts do not: 24 syevd of
// n = 1024 take 271 ms on one stream, 212 ms on four ordinary streams, 94 ms on four green contexts
// of 32 SMs, 67 ms on eight of 16 SMs (tools/try_green_ctx.py).  Events of the primary context work
// on these streams.  Returns how many streams were created (0: not available, use ordinary streams).
int create_green_streams(int device, int num_sms, int want, cudaStream_t* out) {
  if (getenv("KFB_SYEVD_GREEN") && getenv("KFB_SYEVD_GREEN")[0] == '0') return 0;
  typedef CUresult (*GetDev)(CUdevice*, int);
  typedef CUresult (*GetRes)(CUdevice, CUdevResource*, CUdevResourceType);
  typedef CUresult (*Split)(CUdevResource*, unsigned int*, const CUdevResource*, CUdevResource*, unsigned int, unsigned int);
  typedef CUresult (*GenDesc)(CUdevResourceDesc*, CUdevResource*, unsigned int);
  typedef CUresult (*GreenCreate)(CUgreenCtx*, CUdevResourceDesc, CUdevice, unsigned int);
  typedef CUresult (*GreenStream)(CUstream*, CUgreenCtx, unsigned int, int);
  void* fn[6] = {};
  const char* names[6] = {"cuDeviceGet", "cuDeviceGetDevResource", "cuDevSmResourceSplitByCount",
                          "cuDevResourceGenerateDesc", "cuGreenCtxCreate", "cuGreenCtxStreamCreate"};
  for (int i = 0; i < 6; ++i) {
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(names[i], &fn[i], cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
        !fn[i]) {
      cudaGetLastError();
      return 0;
    }
  }
  CUdevice dev;
  CUdevResource sm;
  if (((GetDev)fn[0])(&dev, device) != CUDA_SUCCESS) return 0;
  if (((GetRes)fn[1])(dev, &sm, CU_DEV_RESOURCE_TYPE_SM) != CUDA_SUCCESS) return 0;
  // equal slices, a multiple of 8 SMs each, at most 64
  unsigned int per = (unsigned int)std::min(64, std::max(8, (num_sms / std::max(1, want)) & ~7));
  CUdevResource groups[kMaxLanes], rest;
  unsigned int n = (unsigned int)std::min(want, kMaxLanes);
  if (((Split)fn[2])(groups, &n, &sm, &rest, 0, per) != CUDA_SUCCESS || n == 0) return 0;
  int made = 0;
  for (unsigned int i = 0; i < n; ++i) {
    CUdevResourceDesc desc;
    CUgreenCtx gctx;
    CUstream st;
    if (((GenDesc)fn[3])(&desc, &groups[i], 1) != CUDA_SUCCESS) break;
    if (((GreenCreate)fn[4])(&gctx, desc, dev, CU_GREEN_CTX_DEFAULT_STREAM) != CUDA_SUCCESS) break;
    if (((GreenStream)fn[5])(&st, gctx, CU_STREAM_NON_BLOCKING, 0) != CUDA_SUCCESS) break;
    out[made++] = (cudaStream_t)st;          // the green context lives as long as the process
  }
  return made;
}

int get_pool(kfb::Context* ctx, LanePool** out) {
  std::lock_guard<std::mutex> lock(g_pool_mu);
  LanePool& p = g_pools[ctx->device];
  if (!p.ready) {
    KFB_CUDA(cudaEventCreateWithFlags(&p.start, cudaEventDisableTiming));
    cudaStream_t green[kMaxLanes] = {};
    p.n_green = create_green_streams(ctx->device, ctx->num_sms, lane_count(), green);
    for (int i = 0; i <= kMaxLanes; ++i) {
      Lane& l = i < kMaxLanes ? p.lanes[i] : p.grouped;
      if (i < p.n_green) l.stream = green[i];
      else
      KFB_CUDA(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
      KFB_CUDA(cudaEventCreateWithFlags(&l.done, cudaEventDisableTiming));
      if (cusolverDnCreate(&l.handle) != CUSOLVER_STATUS_SUCCESS)
        KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnCreate failed for worker lane %d", i);
    }
    if (cusolverDnCreateParams(&p.params) != CUSOLVER_STATUS_SUCCESS)
      KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnCreateParams failed");
    p.ready = true;
  }
  *out = &p;
  return KFB_OK;
}

size_t align256(size_t x) { return (x + 255) / 256 * 256; }

size_t lwork_bytes(cusolverDnHandle_t h, int d) {
  int lwork = 0;
  if (cusolverDnSsyevd_bufferSize(h, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, d, nullptr, (d + 3) / 4 * 4, nullptr,
                                  &lwork) != CUSOLVER_STATUS_SUCCESS)
    return 0;
  return align256((size_t)lwork * sizeof(float)) + 256;
}

struct Group {
  int d;
  std::vector<int> idx;
  size_t dev_ws = 0, host_ws = 0;
  size_t off_A = 0, off_W = 0, off_info = 0, off_ws = 0, bytes = 0;   // inside the group region
};
struct BatchPlan {
  std::vector<Group> groups;
  std::vector<std::vector<int>> lanes;   // job indices per worker lane
  std::vector<size_t> lane_off;          // workspace offset per lane (+ end)
  size_t group_off = 0, total = 0;
};

// Which same-size groups go to cusolverDnXsyevBatched.  Measured on B200 (tools/sweep_syev_batched.py,
// tools/bench_syevd.py): the batched solver's time is almost flat in the batch count (it fills the GPU
// whatever the count: b0 + b1*count), syevd's is ~linear in n and per matrix, and the two do not
// overlap well (the batched grids hold the SMs the lanes' microsecond kernels wait for), so the batch
// total is close to  sum(batched groups) + sum(lane work) / 1.33.  A group is batched when that is
// cheaper than leaving its matrices on the lanes.  ResNet-50 on one GPU: 64..512 and 1024 (19 of them)
// batched, 576 x3, 1152 x4, 2048 x7, 2304 x6, 4608 x3 on the lanes; on eight GPUs the per-rank groups
// are small and e.g. two 512s (16 ms batched, 8 ms on a lane) stay on the lanes.
double interp_loglog(int d, const double* tab) {
  static const double kD[10] = {64, 128, 256, 512, 576, 1024, 1152, 2048, 2304, 4608};
  if (d <= kD[0]) return tab[0] * d / kD[0];
  for (int i = 1; i < 10; ++i)
    if (d <= kD[i]) {
      const double t = (std::log((double)d) - std::log(kD[i - 1])) / (std::log(kD[i]) - std::log(kD[i - 1]));
      return std::exp(std::log(tab[i - 1]) + t * (std::log(tab[i]) - std::log(tab[i - 1])));
    }
  return tab[9] * std::pow(d / kD[9], 2.0);
}
bool batched_wins(int d, size_t count, double lane_speedup) {
  static const double kSingle[10] = {0.82, 1.92, 2.17, 5.3, 6.4, 11.3, 13.6, 25.0, 31.0, 88.7};        // syevd, ms
  static const double kB0[10] = {0.35, 0.70, 3.4, 16.2, 25.6, 44.5, 55.0, 123.3, 143.0, 440.0};         // batched, ms
  static const double kB1[10] = {0.001, 0.003, 0.025, 0.19, 0.37, 1.1, 1.5, 2.5, 4.6, 6.0};             // + per matrix
  static const int big_max = []() { const char* e = getenv("KFB_SYEV_BATCH_MAX_D"); return e ? atoi(e) : 0; }();
  static const int big_min_count = []() { const char* e = getenv("KFB_SYEV_BATCH_MIN_COUNT"); return e ? atoi(e) : 8; }();
  if (count < 2) return false;
  if (big_max > 0) return d <= 512 || (d <= big_max && (int)count >= big_min_count);   // explicit thresholds
  const double batched = interp_loglog(d, kB0) + interp_loglog(d, kB1) * (double)count;
  const double lanes = interp_loglog(d, kSingle) * (double)count / lane_speedup;
  return batched < lanes;
}

int make_plan(kfb::Context* ctx, LanePool* pool, const kfb_syevd_job* jobs, int n, BatchPlan* plan) {
  std::vector<int> order(n);
  for (int i = 0; i < n; ++i) {
    KFB_REQUIRE(jobs[i].factor && jobs[i].E && jobs[i].Q && jobs[i].w && jobs[i].dev_info && jobs[i].d > 0,
                "syevd job %d: null argument", i);
    KFB_REQUIRE(jobs[i].ld_factor >= jobs[i].d && jobs[i].ld_e >= jobs[i].d && jobs[i].ld_q >= jobs[i].d,
                "syevd job %d: pitch smaller than d", i);
    order[i] = i;
  }
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return jobs[a].d > jobs[b].d; });
  std::vector<int> singles;
  for (size_t i = 0; i < order.size();) {
    size_t j = i;
    while (j < order.size() && jobs[order[j]].d == jobs[order[i]].d) ++j;
    const int d = jobs[order[i]].d;
    // measured aggregate speed-up of the lanes: 1.33 on ordinary streams, 2.9 / 4.0 with 4 / 8 green contexts
    const double lane_speedup = pool->n_green >= 8 ? 4.0 : pool->n_green >= 4 ? 2.9 : pool->n_green >= 3 ? 2.2 : pool->n_green >= 2 ? 1.6 : 1.33;
    if (batched_wins(d, j - i, lane_speedup)) {
      Group g;
      g.d = d;
      g.idx.assign(order.begin() + i, order.begin() + j);
      const int64_t cnt = (int64_t)g.idx.size();
      if (cusolverDnXsyevBatched_bufferSize(pool->grouped.handle, pool->params, CUSOLVER_EIG_MODE_VECTOR,
                                            CUBLAS_FILL_MODE_UPPER, d, CUDA_R_32F, nullptr, d, CUDA_R_32F, nullptr,
                                            CUDA_R_32F, &g.dev_ws, &g.host_ws, cnt) != CUSOLVER_STATUS_SUCCESS)
        KFB_FAIL(KFB_ERR_CUSOLVER, "XsyevBatched bufferSize failed for d=%d batch=%lld", d, (long long)cnt);
      size_t off = 0;
      g.off_A = off;    off = align256(off + (size_t)cnt * d * d * sizeof(float));
      g.off_W = off;    off = align256(off + (size_t)cnt * d * sizeof(float));
      g.off_info = off; off = align256(off + (size_t)cnt * sizeof(int));
      g.off_ws = off;   off = align256(off + g.dev_ws);
      g.bytes = off;
      plan->groups.push_back(g);
    } else {
      singles.insert(singles.end(), order.begin() + i, order.begin() + j);
    }
    i = j;
  }
  // longest-processing-time placement of the single solves on the lanes (descending d already)
  // without green contexts more than two lanes add nothing (one context's launch path serialises)
  const int lane_cap = pool->n_green > 0 ? pool->n_green : std::min(lane_count(), getenv("KFB_SYEVD_LANES") ? kMaxLanes : 2);
  const int n_lanes = std::min<int>(lane_cap, std::max<size_t>(1, singles.size()));
  plan->lanes.assign(n_lanes, {});
  std::vector<double> load(n_lanes, 0.0);
  for (int i : singles) {
    int best = 0;
    for (int l = 1; l < n_lanes; ++l)
      if (load[l] < load[best]) best = l;
    load[best] += (double)jobs[i].d;          // syevd time is ~linear in d on this GPU (launch-bound panels)
    plan->lanes[best].push_back(i);
  }
  plan->lane_off.assign(n_lanes + 1, 0);
  for (int l = 0; l < n_lanes; ++l) {
    size_t mx = 0;
    for (int i : plan->lanes[l]) mx = std::max(mx, lwork_bytes((cusolverDnHandle_t)ctx->cusolver, jobs[i].d));
    plan->lane_off[l + 1] = plan->lane_off[l] + mx;
  }
  plan->group_off = plan->lane_off[n_lanes];
  size_t gmax = 0;
  for (auto& g : plan->groups) gmax = std::max(gmax, g.bytes);   // groups run one after another
  plan->total = plan->group_off + gmax;
  return KFB_OK;
}

// W (ascending, may be slightly negative) -> w = max(W, 0)
__global__ void clip_copy_kernel(const float* __restrict__ src, float* __restrict__ dst, int d) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < d) dst[i] = fmaxf(src[i], 0.f);
}

int run_group(LanePool* pool, const Group& g, const kfb_syevd_job* jobs, char* region) {
  cudaStream_t st = pool->grouped.stream;
  cusolverDnHandle_t h = pool->grouped.handle;
  if (cusolverDnSetStream(h, st) != CUSOLVER_STATUS_SUCCESS) KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnSetStream failed");
  const int d = g.d;
  const int64_t cnt = (int64_t)g.idx.size();
  float* A = (float*)(region + g.off_A);
  float* W = (float*)(region + g.off_W);
  int* info = (int*)(region + g.off_info);
  for (int64_t b = 0; b < cnt; ++b) {
    const kfb_syevd_job& j = jobs[g.idx[b]];
    KFB_CUDA(cudaMemcpy2DAsync(A + b * d * d, (size_t)d * sizeof(float), j.factor, j.ld_factor * sizeof(float),
                               (size_t)d * sizeof(float), d, cudaMemcpyDeviceToDevice, st));
  }
  std::vector<char> host(std::max<size_t>(g.host_ws, 1));
  cusolverStatus_t rc = cusolverDnXsyevBatched(h, pool->params, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, d,
                                               CUDA_R_32F, A, d, CUDA_R_32F, W, CUDA_R_32F, region + g.off_ws,
                                               g.dev_ws, host.data(), g.host_ws, info, cnt);
  if (rc != CUSOLVER_STATUS_SUCCESS) KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnXsyevBatched failed (%d) for d=%d", (int)rc, d);
  for (int64_t b = 0; b < cnt; ++b) {
    const kfb_syevd_job& j = jobs[g.idx[b]];
    // column-major eigenvectors = rows of the row-major view: that is E; Q is its transpose
    KFB_CUDA(cudaMemcpy2DAsync(j.E, j.ld_e * sizeof(float), A + b * d * d, (size_t)d * sizeof(float),
                               (size_t)d * sizeof(float), d, cudaMemcpyDeviceToDevice, st));
    clip_copy_kernel<<<kfb::ceil_div(d, 256), 256, 0, st>>>(W + b * d, j.w, d);
    KFB_CHECK_LAUNCH();
    int r = kfb::launch_transpose(A + b * d * d, d, d, d, j.Q, j.ld_q, st);
    if (r) return r;
    KFB_CUDA(cudaMemcpyAsync(j.dev_info, info + b, sizeof(int), cudaMemcpyDeviceToDevice, st));
  }
  return KFB_OK;
}
}  // namespace

size_t kfb_syevd_batch_workspace_bytes(const kfb_syevd_job* jobs, int32_t n_jobs) {
  kfb::Context* ctx = kfb::current_context();
  if (!ctx || !jobs || n_jobs <= 0) return 0;
  LanePool* pool;
  if (get_pool(ctx, &pool)) return 0;
  BatchPlan plan;
  if (make_plan(ctx, pool, jobs, n_jobs, &plan)) return 0;
  return plan.total;
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
  BatchPlan plan;
  if ((rc = make_plan(ctx, pool, jobs, n_jobs, &plan))) return rc;
  if (!workspace || workspace_bytes < plan.total)
    KFB_FAIL(KFB_ERR_WORKSPACE, "syevd batch workspace: need %zu bytes, got %zu", plan.total, workspace_bytes);
  KFB_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  const int n_lanes = (int)plan.lanes.size();
  // fork: every lane starts after what is already queued on the caller's stream
  KFB_CUDA(cudaEventRecord(pool->start, stream));
  std::vector<int> status(n_lanes, KFB_OK);
  std::vector<std::string> errors(n_lanes);
  std::vector<std::thread> threads;
  const int device = ctx->device;
  for (int l = 0; l < n_lanes; ++l) {
    if (plan.lanes[l].empty()) continue;
    threads.emplace_back([&, l]() {
      // syevd's host-side panel loop is what takes the time: one thread per lane lets them overlap
      cudaSetDevice(device);
      kfb_init(device);
      Lane& lane = pool->lanes[l];
      if (cudaStreamWaitEvent(lane.stream, pool->start, 0) != cudaSuccess) { status[l] = KFB_ERR_CUDA; return; }
      for (int i : plan.lanes[l]) {
        const kfb_syevd_job& j = jobs[i];
        int r = syevd_one(lane.handle, j.factor, j.ld_factor, j.d, j.E, j.ld_e, j.Q, j.ld_q, j.w,
                          (char*)workspace + plan.lane_off[l], plan.lane_off[l + 1] - plan.lane_off[l], j.dev_info,
                          lane.stream);
        if (r) { status[l] = r; errors[l] = kfb::get_error(); break; }
      }
      if (cudaEventRecord(lane.done, lane.stream) != cudaSuccess) status[l] = KFB_ERR_CUDA;
    });
  }
  // meanwhile, on this thread: the same-size groups through the batched solver
  int grc = KFB_OK;
  if (!plan.groups.empty()) {
    grc = cudaStreamWaitEvent(pool->grouped.stream, pool->start, 0) == cudaSuccess ? KFB_OK : KFB_ERR_CUDA;
    for (size_t gi = 0; gi < plan.groups.size() && grc == KFB_OK; ++gi)
      grc = run_group(pool, plan.groups[gi], jobs, (char*)workspace + plan.group_off);
    if (grc == KFB_OK && cudaEventRecord(pool->grouped.done, pool->grouped.stream) != cudaSuccess) grc = KFB_ERR_CUDA;
  }
  for (auto& t : threads) t.join();
  // join: the caller's stream continues after every lane and the grouped stream
  for (int l = 0; l < n_lanes; ++l)
    if (!plan.lanes[l].empty()) KFB_CUDA(cudaStreamWaitEvent(stream, pool->lanes[l].done, 0));
  if (!plan.groups.empty() && grc == KFB_OK) KFB_CUDA(cudaStreamWaitEvent(stream, pool->grouped.done, 0));
  if (grc) return grc;
  for (int l = 0; l < n_lanes; ++l)
    if (status[l]) KFB_FAIL(status[l], "syevd batch lane %d failed: %s", l, errors[l].c_str());
  return KFB_OK;
}

// ---- explicit damped inverse ------------------------------------------------------------------------
namespace {
__global__ void add_diag_kernel(float* a, int64_t ld, int d, float v) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < d) a[(int64_t)i * ld + i] += v;
}
// potri leaves only one triangle valid: copy it over the other (row-major upper <- lower)
__global__ void __launch_bounds__(256) mirror_kernel(float* a, int64_t ld, int d) {
  const int i = blockIdx.y * 16 + threadIdx.y, j = blockIdx.x * 16 + threadIdx.x;
  if (i < d && j < d && j > i) a[(int64_t)i * ld + j] = a[(int64_t)j * ld + i];
}
}  // namespace

size_t kfb_cholesky_inverse_workspace_bytes(int32_t d) {
  kfb::Context* ctx = kfb::current_context();
  if (!ctx || d <= 0) return 0;
  int l1 = 0, l2 = 0;
  cusolverDnHandle_t h = (cusolverDnHandle_t)ctx->cusolver;
  if (cusolverDnSpotrf_bufferSize(h, CUBLAS_FILL_MODE_UPPER, d, nullptr, (d + 3) / 4 * 4, &l1) != CUSOLVER_STATUS_SUCCESS) return 0;
  if (cusolverDnSpotri_bufferSize(h, CUBLAS_FILL_MODE_UPPER, d, nullptr, (d + 3) / 4 * 4, &l2) != CUSOLVER_STATUS_SUCCESS) return 0;
  return (size_t)std::max(l1, l2) * sizeof(float) + 256;
}

int kfb_cholesky_inverse(const float* factor, int64_t ld_factor, int32_t d, float damping, float* out,
                         int64_t ld_out, void* workspace, size_t workspace_bytes, int32_t* dev_info, void* stream_) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(factor && out && dev_info && d > 0 && ld_factor >= d && ld_out >= d, "bad cholesky_inverse arguments");
  cudaStream_t stream = (cudaStream_t)stream_;
  cusolverDnHandle_t h = (cusolverDnHandle_t)ctx->cusolver;
  if (cusolverDnSetStream(h, stream) != CUSOLVER_STATUS_SUCCESS) KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnSetStream failed");
  int l1 = 0, l2 = 0;
  if (cusolverDnSpotrf_bufferSize(h, CUBLAS_FILL_MODE_UPPER, d, out, (int)ld_out, &l1) != CUSOLVER_STATUS_SUCCESS ||
      cusolverDnSpotri_bufferSize(h, CUBLAS_FILL_MODE_UPPER, d, out, (int)ld_out, &l2) != CUSOLVER_STATUS_SUCCESS)
    KFB_FAIL(KFB_ERR_CUSOLVER, "potrf/potri bufferSize failed for d=%d", d);
  const size_t need = (size_t)std::max(l1, l2) * sizeof(float);
  if (!workspace || workspace_bytes < need)
    KFB_FAIL(KFB_ERR_WORKSPACE, "cholesky_inverse workspace: need %zu bytes, got %zu", need, workspace_bytes);
  KFB_CUDA(cudaMemcpy2DAsync(out, ld_out * sizeof(float), factor, ld_factor * sizeof(float), (size_t)d * sizeof(float), d,
                             cudaMemcpyDeviceToDevice, stream));
  add_diag_kernel<<<kfb::ceil_div(d, 256), 256, 0, stream>>>(out, ld_out, d, damping);
  KFB_CHECK_LAUNCH();
  // the symmetric matrix read column-major is itself; cuSOLVER's UPPER triangle (column-major) is
  // the LOWER triangle of the row-major view, which mirror_kernel copies to the upper one
  if (cusolverDnSpotrf(h, CUBLAS_FILL_MODE_UPPER, d, out, (int)ld_out, (float*)workspace, l1, dev_info) != CUSOLVER_STATUS_SUCCESS)
    KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnSpotrf failed for d=%d", d);
  if (cusolverDnSpotri(h, CUBLAS_FILL_MODE_UPPER, d, out, (int)ld_out, (float*)workspace, l2, dev_info) != CUSOLVER_STATUS_SUCCESS)
    KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnSpotri failed for d=%d", d);
  dim3 grid(kfb::ceil_div(d, 16), kfb::ceil_div(d, 16)), block(16, 16);
  mirror_kernel<<<grid, block, 0, stream>>>(out, ld_out, d);
  KFB_CHECK_LAUNCH();
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
