// Factor SYRK on the 5th-generation tensor cores:  F += w * X^T X  for one 128x128 upper-triangle
// tile pair and one K chunk per work item, X read IN PLACE from the captured NCHW tensors
// (implicit im2col: the unfolded matrix of kfac/layers/conv.py:50-70 is never materialised).
//
//   warps 0-3   epilogue: tcgen05.ld -> x weight -> red.global.add.v4.f32 into the job's scratch
//   warp  4     MMA issuer: tcgen05.mma.kind::tf32, M=128 N=128 K=8, fp32 accumulators in TMEM
//               (double buffered); a diagonal tile pair uses ONE smem tile for both operands
//   warps 5-12  producers: gather X^T tiles (128 features x 32 rows of X, K-major) from global
//               memory straight into the 128B-swizzled UMMA shared-memory layout
//               - NCHW operands with 16-byte alignable rows: 128-bit loads/stores
//               - im2col operands: one 32-bit load per element with a per-lane validity bitmask
//                 over (ky,kx) (padding / image border), coalesced along the output row
// Operands stay fp32 in shared memory; kind::tf32 reads the top 19 bits, so no conversion pass.
#include <algorithm>
#include <vector>

#include "kfb_factor.cuh"
#include "kfb_umma.cuh"

namespace kfb {
namespace {

constexpr int TM = 128;                 // tile side (features)
constexpr int BK = 32;                  // rows of X per stage = one 128-byte swizzle row
constexpr int UMMA_K = 8;
constexpr int kStages = 5;
constexpr int kTileBytes = TM * BK * 4; // 16 KB
constexpr int kStageBytes = 2 * kTileBytes;
constexpr int kProducerWarps = 8;
constexpr int kThreads = 32 * (4 + 1 + kProducerWarps);
constexpr int kTmemCols = 256;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;

constexpr int kBitNever = 62;           // row beyond d: always zero
constexpr int kBitBias = 61;            // bias feature: 1 wherever the X row exists
constexpr int kBitAlways = 63;          // NCHW: valid wherever the X row exists

struct RowInfo { int off; int bit; };

__device__ __forceinline__ RowInfo row_info(const DevSegment& s, int d, int i) {
  RowInfo r;
  r.off = 0;
  if (i >= d) { r.bit = kBitNever; return r; }
  if (s.has_bias && i == d - 1) { r.bit = kBitBias; return r; }
  if (s.kind == KFB_SEG_NCHW) {
    r.off = i * s.h * s.w;
    r.bit = kBitAlways;
  } else {
    int khw = s.kh * s.kw;
    int ci = i / khw;
    int q = i - ci * khw;
    int ky = q / s.kw, kx = q - ky * s.kw;
    r.off = (ci * s.h + ky) * s.w + kx;
    r.bit = q;
  }
  return r;
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(kThreads, 1) syrk_tc_kernel(const DevJob* __restrict__ jobs,
                                                            const DevSegment* __restrict__ segs,
                                                            const TcItem* __restrict__ items, int n_items) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + kStages * kStageBytes);
  uint64_t* empty = full + kStages;
  uint64_t* tfull = empty + kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = (uint32_t*)(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { umma::mbar_init(&full[s], kProducerWarps); umma::mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { umma::mbar_init(&tfull[s], 1); umma::mbar_init(&tempty[s], 128); }
    umma::fence_barrier_init();
  }
  if (warp == 4) umma::tmem_alloc<kTmemCols>(tmem_slot);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= 5) {
    // ===================== producers =====================
    const int pw = warp - 5;
    int stage = 0; uint32_t phase = 0;
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
      const TcItem item = items[it];
      const DevJob& job = jobs[item.job];
      const DevSegment s = segs[item.seg];
      const int d = job.d;
      const bool diag = item.ti == item.tj;
      // lane j holds the decode of tile row (pw + 8*j'): j' = j & 15, tile A for j < 16, B otherwise
      const int my_row = pw + kProducerWarps * (lane & 15);
      const RowInfo mine = row_info(s, d, (lane < 16 ? item.ti : item.tj) * TM + my_row);
      const int S = s.oh * s.ow;
      const bool vec4 = s.kind == KFB_SEG_NCHW && (S & 3) == 0 && (((uintptr_t)s.data) & 15) == 0;
      const int n_tiles = diag ? 1 : 2;

      for (int64_t k0 = item.k_begin; k0 < item.k_end; k0 += BK) {
        umma::mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* st = smem + stage * kStageBytes;
        if (vec4) {
          // 8 lanes x float4 cover one tile row (32 k); 4 rows per warp instruction
          const int kq = lane & 7, rsub = lane >> 3;
          const int64_t k = k0 + 4 * kq;
          const bool kin = k < item.k_end;
          int64_t base = 0;
          if (kin) {
            int64_t n = k / S;
            base = n * (int64_t)s.c * S + (k - n * S);
          }
          for (int t = 0; t < n_tiles; ++t) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int j = 4 * i + rsub;                       // which of this warp's 16 rows
              const int off = __shfl_sync(0xffffffffu, mine.off, j + 16 * t);
              const int bit = __shfl_sync(0xffffffffu, mine.bit, j + 16 * t);
              float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
              if (kin && bit == kBitAlways) v = __ldg(reinterpret_cast<const float4*>(s.data + base + off));
              const int r = pw + kProducerWarps * j;
              *reinterpret_cast<float4*>(st + t * kTileBytes + r * 128 + ((kq ^ (r & 7)) << 4)) = v;
            }
          }
        } else {
          const int64_t k = k0 + lane;
          const bool kin = k < item.k_end;
          int64_t base = 0;
          uint64_t mask = 0;
          if (kin) {
            int64_t n = k / S;
            int r = (int)(k - n * S);
            if (s.kind == KFB_SEG_NCHW) {
              base = n * (int64_t)s.c * S + r;
              mask = (1ull << kBitAlways) | (1ull << kBitBias);
            } else {
              int oy = r / s.ow, ox = r - oy * s.ow;
              int y0 = oy * s.sh - s.ph, x0 = ox * s.sw - s.pw;
              base = n * (int64_t)s.c * s.h * s.w + (int64_t)y0 * s.w + x0;
              int kx_lo = max(0, -x0), kx_hi = min(s.kw, s.w - x0);
              int ky_lo = max(0, -y0), ky_hi = min(s.kh, s.h - y0);
              uint64_t xm = kx_hi > kx_lo ? (((1ull << (kx_hi - kx_lo)) - 1) << kx_lo) : 0;
              for (int ky = ky_lo; ky < ky_hi; ++ky) mask |= xm << (ky * s.kw);
              mask |= 1ull << kBitBias;
            }
          }
          for (int t = 0; t < n_tiles; ++t) {
#pragma unroll 4
            for (int j = 0; j < 16; ++j) {
              const int off = __shfl_sync(0xffffffffu, mine.off, j + 16 * t);
              const int bit = __shfl_sync(0xffffffffu, mine.bit, j + 16 * t);
              float v = 0.f;
              if ((mask >> bit) & 1ull) v = (bit == kBitBias) ? 1.f : __ldg(s.data + base + off);
              const int r = pw + kProducerWarps * j;
              *reinterpret_cast<float*>(st + t * kTileBytes + r * 128 + (((lane >> 2) ^ (r & 7)) << 4) +
                                        ((lane & 3) << 2)) = v;
            }
          }
        }
        umma::fence_proxy_async_smem();      // generic-proxy stores -> visible to the tensor core
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(&full[stage]);
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 4) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma::make_idesc_tf32(TM, TM);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const TcItem item = items[it];
        const bool diag = item.ti == item.tj;
        umma::mbar_wait(&tempty[acc], acc_phase ^ 1);
        umma::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * TM;
        bool first = true;
        for (int64_t k0 = item.k_begin; k0 < item.k_end; k0 += BK) {
          umma::mbar_wait(&full[stage], phase);
          umma::tc_fence_after();
          const uint32_t sa = umma::smem_u32(smem + stage * kStageBytes);
          const uint64_t da = umma::make_kmajor_sw128_desc(sa);
          const uint64_t db = diag ? da : umma::make_kmajor_sw128_desc(sa + kTileBytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            umma::mma_tf32_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (first && k == 0) ? 0u : 1u);
          }
          first = false;
          umma::tc_commit(&empty[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma::tc_commit(&tfull[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 0..3) =====================
    const int q = warp;
    int acc = 0; uint32_t acc_phase = 0;
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
      const TcItem item = items[it];
      const DevJob& job = jobs[item.job];
      const float weight = segs[item.seg].weight;
      const int d = job.d;
      const int i = item.ti * TM + q * 32 + lane;
      const int j0 = item.tj * TM;
      umma::mbar_wait(&tfull[acc], acc_phase);
      umma::tc_fence_after();
      const uint32_t taddr = tmem_base + acc * TM + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < TM / 32; ++c) {
        uint32_t r[32];
        umma::tmem_ld_32x32(taddr + c * 32, r);
        umma::tmem_ld_wait();
        const int jb = j0 + c * 32;
        if (i < d && jb < d && jb + 31 >= i) {
          float* row = job.scratch + (int64_t)i * job.ld_scratch + jb;
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const int j = jb + 4 * g;
            if (j + 3 < d && j >= i) {
              red_add_v4(row + 4 * g, __uint_as_float(r[4 * g]) * weight, __uint_as_float(r[4 * g + 1]) * weight,
                         __uint_as_float(r[4 * g + 2]) * weight, __uint_as_float(r[4 * g + 3]) * weight);
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (j + e < d && j + e >= i) atomicAdd(row + 4 * g + e, __uint_as_float(r[4 * g + e]) * weight);
            }
          }
        }
      }
      umma::tc_fence_before();
      umma::mbar_arrive(&tempty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  umma::tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    umma::tc_fence_after();
    umma::tmem_dealloc<kTmemCols>(tmem_base);
  }
}

}  // namespace

// Build the tensor-core work list for the jobs marked use_tc: (tile pair, K chunk) items, longest first.
void plan_syrk_tc(FactorPlan* plan) {
  constexpr int64_t kChunk = 4096;
  plan->tc_items.clear();
  for (size_t ji = 0; ji < plan->jobs.size(); ++ji) {
    const DevJob& j = plan->jobs[ji];
    if (!j.use_tc) continue;
    const int T = ceil_div(j.d, TM);
    for (int si = 0; si < j.n_segments; ++si) {
      const DevSegment& s = plan->segs[j.seg_start + si];
      const int64_t nch = std::max<int64_t>(1, ceil_div64(s.K, kChunk));
      const int64_t len = ceil_div64(ceil_div64(s.K, nch), BK) * BK;
      for (int64_t k = 0; k < s.K; k += len)
        for (int ti = 0; ti < T; ++ti)
          for (int tj = ti; tj < T; ++tj) {
            TcItem it;
            it.job = (int)ji; it.seg = j.seg_start + si; it.ti = ti; it.tj = tj;
            it.k_begin = k; it.k_end = std::min<int64_t>(s.K, k + len);
            plan->tc_items.push_back(it);
          }
    }
  }
  std::stable_sort(plan->tc_items.begin(), plan->tc_items.end(), [](const TcItem& a, const TcItem& b) {
    int64_t ca = (a.k_end - a.k_begin) * (a.ti == a.tj ? 1 : 2), cb = (b.k_end - b.k_begin) * (b.ti == b.tj ? 1 : 2);
    return ca > cb;
  });
}

int launch_syrk_tc(const FactorPlan& plan, void* workspace, Context* ctx, cudaStream_t stream) {
  if (plan.tc_items.empty()) return KFB_OK;
  static bool attr_set = false;
  if (!attr_set) {
    KFB_CUDA(cudaFuncSetAttribute(syrk_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  char* ws = (char*)workspace;
  const int n = (int)plan.tc_items.size();
  const int grid = std::min(n, ctx->num_sms);
  syrk_tc_kernel<<<grid, kThreads, kSmemBytes, stream>>>((const DevJob*)(ws + plan.off_jobs),
                                                       (const DevSegment*)(ws + plan.off_segs),
                                                       (const TcItem*)(ws + plan.off_tc_items), n);
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

int launch_syrk(const FactorPlan& plan, void* workspace, Context* ctx, cudaStream_t stream) {
  int rc = launch_syrk_tc(plan, workspace, ctx, stream);
  if (rc) return rc;
  return launch_syrk_simt(plan, workspace, stream);
}

}  // namespace kfb
