// AMP support around the factor kernels: the on-device capture filter (which output-gradient captures
// are finite, the unscale factor, the renormalisation when captures are dropped) and the 16-bit
// narrow / widen passes for inv_dtype = float16 / bfloat16.  All HBM-bound elementwise work.
#include <algorithm>
#include <mutex>

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "kfb_common.cuh"

namespace kfb {
namespace {

struct FilterDev {
  const void* data;
  int64_t numel;
  int dtype;
  int group;
  float rows;
  float host_inv_scale_sq;
  const float* loss_scale;
  float* dev_scale;
  int first_block;       // blocks [first_block, next first_block) scan this capture
};
struct GroupDev {
  float* dev_rescale;
  int renormalise;
  int first_job, n_jobs; // captures of this group are contiguous in the job array
};

__device__ __forceinline__ bool finite_bits32(uint32_t u) { return (u & 0x7f800000u) != 0x7f800000u; }
__device__ __forceinline__ bool finite_f16(unsigned short u) { return (u & 0x7c00u) != 0x7c00u; }
__device__ __forceinline__ bool finite_bf16(unsigned short u) { return (u & 0x7f80u) != 0x7f80u; }

// Descriptors travel as a by-value kernel parameter (no device allocation, no upload).
constexpr int kFilterBatch = 192;
struct FilterBatch {
  int n_jobs, n_groups;
  int first[kFilterBatch + 1];
  FilterDev jobs[kFilterBatch];
  GroupDev groups[kFilterBatch];
};

// pass 0: dev_scale[c] = 1 / s^2 (every capture starts as "kept")
__global__ void filter_init_kernel(const __grid_constant__ FilterBatch batch) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= batch.n_jobs) return;
  const FilterDev& j = batch.jobs[i];
  float inv = j.host_inv_scale_sq;
  if (j.loss_scale) {
    const float s = __ldg(j.loss_scale);
    inv = 1.f / (s * s);
  }
  *j.dev_scale = inv;
}

// pass 1: any inf / nan in a capture zeroes its dev_scale.  16 bytes per thread and step; a block that has
// seen a non-finite value stops early.
__global__ void __launch_bounds__(256) filter_scan_kernel(const __grid_constant__ FilterBatch batch) {
  const FilterDev* jobs = batch.jobs;
  const int* first = batch.first;
  int lo = 0, hi = batch.n_jobs;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (first[mid] <= (int)blockIdx.x) lo = mid; else hi = mid;
  }
  const FilterDev& j = jobs[lo];
  const int nblocks = first[lo + 1] - first[lo];
  const int b = blockIdx.x - first[lo];
  const int per16 = j.dtype == KFB_F32 ? 4 : 8;                 // elements per 16 bytes
  const int64_t vecs = ((uintptr_t)j.data & 15) == 0 ? j.numel / per16 : 0;
  bool bad = false;
  const uint4* v = reinterpret_cast<const uint4*>(j.data);
  for (int64_t i = (int64_t)b * 256 + threadIdx.x; i < vecs && !bad; i += (int64_t)nblocks * 256) {
    const uint4 q = __ldg(v + i);
    if (j.dtype == KFB_F32) {
      bad = !(finite_bits32(q.x) && finite_bits32(q.y) && finite_bits32(q.z) && finite_bits32(q.w));
    } else {
      const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const unsigned short a = (unsigned short)(w[k] & 0xffff), c = (unsigned short)(w[k] >> 16);
        bad |= j.dtype == KFB_F16 ? !(finite_f16(a) && finite_f16(c)) : !(finite_bf16(a) && finite_bf16(c));
      }
    }
  }
  // tail (and everything, for a misaligned capture): scalar
  for (int64_t i = vecs * per16 + (int64_t)b * 256 + threadIdx.x; i < j.numel && !bad; i += (int64_t)nblocks * 256) {
    if (j.dtype == KFB_F32) bad = !finite_bits32(__float_as_uint(__ldg(reinterpret_cast<const float*>(j.data) + i)));
    else {
      const unsigned short u = __ldg(reinterpret_cast<const unsigned short*>(j.data) + i);
      bad = j.dtype == KFB_F16 ? !finite_f16(u) : !finite_bf16(u);
    }
  }
  if (__syncthreads_or(bad) && threadIdx.x == 0) *j.dev_scale = 0.f;
}

// pass 2: per factor, what is left of its rows
__global__ void filter_group_kernel(const __grid_constant__ FilterBatch batch) {
  const FilterDev* jobs = batch.jobs;
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= batch.n_groups) return;
  const GroupDev& gr = batch.groups[g];
  float all = 0.f, kept = 0.f;
  for (int i = 0; i < gr.n_jobs; ++i) {
    const FilterDev& j = jobs[gr.first_job + i];
    all += j.rows;
    if (*j.dev_scale != 0.f) kept += j.rows;
  }
  *gr.dev_rescale = kept == 0.f ? -1.f : (gr.renormalise ? all / kept : 1.f);
}

template <int kDtype>
__global__ void __launch_bounds__(256) narrow_kernel(const float* __restrict__ src, unsigned short* __restrict__ dst, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i >= n) return;
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  if (i + 3 < n && (((uintptr_t)(src + i)) & 15) == 0) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(src + i));
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  } else {
    for (int k = 0; k < 4 && i + k < n; ++k) v[k] = src[i + k];
  }
  unsigned short o[4];
#pragma unroll
  for (int k = 0; k < 4; ++k)
    o[k] = kDtype == KFB_F16 ? __half_as_ushort(__float2half_rn(v[k])) : __bfloat16_as_ushort(__float2bfloat16_rn(v[k]));
  if (i + 3 < n && (((uintptr_t)(dst + i)) & 7) == 0) {
    *reinterpret_cast<uint2*>(dst + i) = make_uint2((uint32_t)o[0] | ((uint32_t)o[1] << 16), (uint32_t)o[2] | ((uint32_t)o[3] << 16));
  } else {
    for (int k = 0; k < 4 && i + k < n; ++k) dst[i + k] = o[k];
  }
}

template <int kDtype>
__global__ void __launch_bounds__(256) widen_kernel(const unsigned short* __restrict__ src, float* __restrict__ dst, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i >= n) return;
  unsigned short u[4] = {0, 0, 0, 0};
  if (i + 3 < n && (((uintptr_t)(src + i)) & 7) == 0) {
    const uint2 q = __ldg(reinterpret_cast<const uint2*>(src + i));
    u[0] = (unsigned short)(q.x & 0xffff); u[1] = (unsigned short)(q.x >> 16);
    u[2] = (unsigned short)(q.y & 0xffff); u[3] = (unsigned short)(q.y >> 16);
  } else {
    for (int k = 0; k < 4 && i + k < n; ++k) u[k] = src[i + k];
  }
  float v[4];
#pragma unroll
  for (int k = 0; k < 4; ++k)
    v[k] = kDtype == KFB_F16 ? __half2float(__ushort_as_half(u[k])) : __uint_as_float((uint32_t)u[k] << 16);
  if (i + 3 < n && (((uintptr_t)(dst + i)) & 15) == 0) {
    *reinterpret_cast<float4*>(dst + i) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
    for (int k = 0; k < 4 && i + k < n; ++k) dst[i + k] = v[k];
  }
}

}  // namespace
}  // namespace kfb

extern "C" {

int kfb_capture_filter(const kfb_filter_job* jobs, int32_t n_jobs, const kfb_filter_group* groups, int32_t n_groups,
                       void* stream_) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(n_jobs >= 0 && n_groups > 0 && (jobs || n_jobs == 0) && groups, "bad filter arguments");
  if (n_jobs == 0) return KFB_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  for (int g = 0; g < n_groups; ++g) KFB_REQUIRE(groups[g].dev_rescale, "filter group %d: null output", g);
  static kfb::FilterBatch batch;        // 20 KB: kept off the stack; calls are serialised
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  int i = 0;
  while (i < n_jobs) {
    // one launch set per run of whole groups that fits a batch
    batch.n_jobs = 0;
    batch.n_groups = 0;
    int blocks = 0;
    while (i < n_jobs) {
      const int g = jobs[i].group;
      KFB_REQUIRE(g >= 0 && g < n_groups, "filter job %d: group %d out of range", i, g);
      int e = i;
      while (e < n_jobs && jobs[e].group == g) ++e;
      KFB_REQUIRE(e - i <= kfb::kFilterBatch, "more than %d captures in one factor", kfb::kFilterBatch);
      if (batch.n_jobs + (e - i) > kfb::kFilterBatch || batch.n_groups == kfb::kFilterBatch) break;
      kfb::GroupDev& gd = batch.groups[batch.n_groups++];
      gd.dev_rescale = groups[g].dev_rescale;
      gd.renormalise = groups[g].renormalise;
      gd.first_job = batch.n_jobs;
      gd.n_jobs = e - i;
      for (; i < e; ++i) {
        const kfb_filter_job& j = jobs[i];
        KFB_REQUIRE(j.data && j.numel > 0 && j.dev_scale, "filter job %d: null argument", i);
        KFB_REQUIRE(j.dtype >= KFB_F32 && j.dtype <= KFB_BF16, "filter job %d: unknown dtype %d", i, j.dtype);
        kfb::FilterDev& d = batch.jobs[batch.n_jobs];
        d.data = j.data; d.numel = j.numel; d.dtype = j.dtype; d.group = batch.n_groups - 1; d.rows = j.rows;
        d.host_inv_scale_sq = j.host_inv_scale_sq; d.loss_scale = j.loss_scale; d.dev_scale = j.dev_scale;
        d.first_block = blocks;
        batch.first[batch.n_jobs++] = blocks;
        // 64 KB per block and pass, at most 4 CTAs per SM for one capture
        const int64_t bytes = j.numel * (j.dtype == KFB_F32 ? 4 : 2);
        blocks += (int)std::max<int64_t>(1, std::min<int64_t>((bytes + 65535) / 65536, 4 * (int64_t)ctx->num_sms));
      }
    }
    batch.first[batch.n_jobs] = blocks;
    kfb::filter_init_kernel<<<kfb::ceil_div(batch.n_jobs, 128), 128, 0, stream>>>(batch);
    KFB_CHECK_LAUNCH();
    kfb::filter_scan_kernel<<<blocks, 256, 0, stream>>>(batch);
    KFB_CHECK_LAUNCH();
    kfb::filter_group_kernel<<<kfb::ceil_div(batch.n_groups, 128), 128, 0, stream>>>(batch);
    KFB_CHECK_LAUNCH();
  }
  return KFB_OK;
}

int kfb_narrow(const float* src, void* dst16, int64_t n, int32_t dtype, void* stream) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(src && dst16 && n > 0 && (dtype == KFB_F16 || dtype == KFB_BF16), "bad narrow arguments");
  const unsigned int blocks = (unsigned int)((n + 1023) / 1024);
  if (dtype == KFB_F16) kfb::narrow_kernel<KFB_F16><<<blocks, 256, 0, (cudaStream_t)stream>>>(src, (unsigned short*)dst16, n);
  else kfb::narrow_kernel<KFB_BF16><<<blocks, 256, 0, (cudaStream_t)stream>>>(src, (unsigned short*)dst16, n);
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

int kfb_widen(const void* src16, float* dst, int64_t n, int32_t dtype, void* stream) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(src16 && dst && n > 0 && (dtype == KFB_F16 || dtype == KFB_BF16), "bad widen arguments");
  const unsigned int blocks = (unsigned int)((n + 1023) / 1024);
  if (dtype == KFB_F16) kfb::widen_kernel<KFB_F16><<<blocks, 256, 0, (cudaStream_t)stream>>>((const unsigned short*)src16, dst, n);
  else kfb::widen_kernel<KFB_BF16><<<blocks, 256, 0, (cudaStream_t)stream>>>((const unsigned short*)src16, dst, n);
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

}  // extern "C"
