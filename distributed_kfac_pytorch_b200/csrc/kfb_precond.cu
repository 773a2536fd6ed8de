// Preconditioning: pack [dW | db], four grouped GEMM stages over all layers, the KL-clip
// inner product, and the scaled in-place write-back.
#include <algorithm>
#include <vector>

#include "kfb_gemm.cuh"

namespace kfb {
namespace {

// ============================================================================================
// SIMT grouped GEMM (exact fp32 FMA): 64x64 tile, K step 32, 4x4 per thread
// ============================================================================================
constexpr int TS = 64, BK = 32, NT = 256;

__global__ void __launch_bounds__(NT) gemm_simt_kernel(const __grid_constant__ GemmBatch batch, float* dot_acc) {
  __shared__ float As[BK][TS + 1];
  __shared__ float Bs[BK][TS + 1];
  __shared__ float red[NT / 32];
  const int jidx = find_job(batch, blockIdx.x);
  const GemmProb p = batch.jobs[jidx];
  const int local = blockIdx.x - batch.start[jidx];
  const int tiles_n = (p.N + TS - 1) / TS;
  const int m0 = (local / tiles_n) * TS, n0 = (local % tiles_n) * TS;
  const int t = threadIdx.x;
  const int kk_l = t & 31, r_l = t >> 5;
  const int tx = t & 15, ty = t >> 4;
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;

  for (int k0 = 0; k0 < p.K; k0 += BK) {
    const int k = k0 + kk_l;
    const bool kin = k < p.K;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int r = r_l + j * 8;
      int m = m0 + r, n = n0 + r;
      As[kk_l][r] = (kin && m < p.M) ? __ldg(p.A + (int64_t)m * p.lda + k) : 0.f;
      Bs[kk_l][r] = (kin && n < p.N) ? __ldg(p.B + (int64_t)n * p.ldb + k) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { a[q] = As[kk][ty * 4 + q]; b[q] = Bs[kk][tx * 4 + q]; }
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int r = 0; r < 4; ++r) acc[q][r] = fmaf(a[q], b[r], acc[q][r]);
    }
    __syncthreads();
  }

  float dot = 0.f;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    int m = m0 + ty * 4 + q;
    if (m >= p.M) continue;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      int n = n0 + tx * 4 + r;
      if (n >= p.N) continue;
      float v = acc[q][r];
      if (p.Mul) v *= p.Mul[(int64_t)m * p.ldmul + n];
      p.D[(int64_t)m * p.ldd + n] = v;
      if (p.Dot) dot = fmaf(v, p.Dot[(int64_t)m * p.lddot + n], dot);
    }
  }
  if (p.Dot && dot_acc) {
    dot = warp_sum(dot);
    if ((t & 31) == 0) red[t >> 5] = dot;
    __syncthreads();
    if (t < 32) {
      float v = t < NT / 32 ? red[t] : 0.f;
      v = warp_sum(v);
      if (t == 0) atomicAdd(dot_acc, v);
    }
  }
}

// ============================================================================================
// gradient pack / dot / write-back (HBM-bound multi-tensor kernels)
// ============================================================================================
struct GradJob {
  float* wg;
  float* bg;
  float* v;        // pack: destination gmat; dot / writeback: source v
  int64_t ld;
  int da, dg, nw;  // nw = da - has_bias
  int pad_;
};
constexpr int kGradBatch = 512;
typedef Batch<GradJob, kGradBatch> GradBatch;
constexpr int kElemsPerBlock = 8192;     // write-back: 32 KB read + 32 KB written per block, eight 16-byte accesses per thread in flight
// pack / dot touch a few small layers (conv1 and fc in a ResNet: 2 M elements): small blocks so that the
// handful of layers still spreads over every SM
template <int kMode> struct GradBlock { static constexpr int kElems = kMode == 2 ? kElemsPerBlock : 2048; };

enum { kPack = 0, kDot = 1, kWrite = 2 };

template <int kMode>
__global__ void __launch_bounds__(256) grad_kernel(const __grid_constant__ GradBatch batch, float* acc_out,
                                                   const float* vg_acc, float lr, float kl_clip, float fixed_scale) {
  __shared__ float red[8];
  const int jidx = find_job(batch, blockIdx.x);
  const GradJob j = batch.jobs[jidx];
  const int64_t total = (int64_t)j.dg * j.da;
  constexpr int kEPB = GradBlock<kMode>::kElems;
  const int64_t e0 = (int64_t)(blockIdx.x - batch.start[jidx]) * kEPB;
  float nu = fixed_scale;
  if (kMode == kWrite && vg_acc) {
    float vg = *vg_acc * lr * lr;
    if (vg != 0.f) nu = fminf(1.f, sqrtf(kl_clip / fabsf(vg)));
  }
  float dot = 0.f;
  // weight-only layers whose matrix rows are unpadded 16-byte multiples (every ResNet conv but the
  // first): matrix and gradient are the same flat array, moved 16 bytes per thread
  if (kMode == kWrite && j.nw == j.da && j.ld == j.da && (j.da & 3) == 0 && (((uintptr_t)j.wg | (uintptr_t)j.v) & 15) == 0) {
    float4 v[kElemsPerBlock / 1024];
#pragma unroll
    for (int u = 0; u < kElemsPerBlock / 1024; ++u) {          // all loads first: eight lines in flight per thread
      const int64_t e = e0 + (int64_t)(u * 256 + threadIdx.x) * 4;
      v[u] = e < total ? __ldcs(reinterpret_cast<const float4*>(j.v + e)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < kElemsPerBlock / 1024; ++u) {
      const int64_t e = e0 + (int64_t)(u * 256 + threadIdx.x) * 4;
      if (e >= total) break;
      *reinterpret_cast<float4*>(j.wg + e) = make_float4(v[u].x * nu, v[u].y * nu, v[u].z * nu, v[u].w * nu);
    }
    return;
  }
  const bool small = total < ((int64_t)1 << 31);          // 32-bit division: a 64-bit one costs ~100 instructions
#pragma unroll
  for (int u = 0; u < kEPB / 256; ++u) {
    int64_t e = e0 + u * 256 + threadIdx.x;
    if (e >= total) break;
    int r = small ? (int)((uint32_t)e / (uint32_t)j.da) : (int)(e / j.da), c = (int)(e - (int64_t)r * j.da);
    float* gp = c < j.nw ? j.wg + (int64_t)r * j.nw + c : j.bg + r;
    float* vp = j.v + (int64_t)r * j.ld + c;
    if (kMode == kPack) *vp = *gp;
    else if (kMode == kDot) dot = fmaf(*vp, *gp, dot);
    else *gp = nu * *vp;
  }
  if (kMode == kDot) {
    dot = warp_sum(dot);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = dot;
    __syncthreads();
    if (threadIdx.x < 32) {
      float v = threadIdx.x < 8 ? red[threadIdx.x] : 0.f;
      v = warp_sum(v);
      if (threadIdx.x == 0) atomicAdd(acc_out, v);
    }
  }
}

template <int kMode>
int run_grad(const GradJob* jobs, int n, float* acc_out, const float* vg_acc, float lr, float kl_clip,
             float fixed_scale, cudaStream_t stream) {
  for (int base = 0; base < n; base += kGradBatch) {
    GradBatch batch;
    batch.n = std::min(kGradBatch, n - base);
    int blocks = 0;
    for (int i = 0; i < batch.n; ++i) {
      batch.jobs[i] = jobs[base + i];
      batch.start[i] = blocks;
      blocks += (int)ceil_div64((int64_t)jobs[base + i].dg * jobs[base + i].da, GradBlock<kMode>::kElems);
    }
    batch.start[batch.n] = blocks;
    grad_kernel<kMode><<<blocks, 256, 0, stream>>>(batch, acc_out, vg_acc, lr, kl_clip, fixed_scale);
    KFB_CHECK_LAUNCH();
  }
  return KFB_OK;
}

int convert_grad_jobs(const kfb_grad_job* jobs, int n, std::vector<GradJob>* out) {
  KFB_REQUIRE(n >= 0 && (jobs || n == 0), "bad job array");
  out->resize(n);
  for (int i = 0; i < n; ++i) {
    const kfb_grad_job& j = jobs[i];
    KFB_REQUIRE(j.weight_grad && j.v && j.da > 0 && j.dg > 0 && j.ld_v >= j.da, "grad job %d is malformed", i);
    KFB_REQUIRE(!j.has_bias || j.bias_grad, "grad job %d: has_bias without bias_grad", i);
    GradJob& g = (*out)[i];
    g.wg = j.weight_grad; g.bg = j.bias_grad; g.v = (float*)j.v; g.ld = j.ld_v;
    g.da = j.da; g.dg = j.dg; g.nw = j.da - (j.has_bias ? 1 : 0); g.pad_ = 0;
  }
  return KFB_OK;
}

}  // namespace

int gemm_grouped_simt(const GemmProb* probs, int n, float* dot_acc, cudaStream_t stream) {
  for (int base = 0; base < n; base += kGemmBatch) {
    GemmBatch batch;
    batch.n = std::min(kGemmBatch, n - base);
    int blocks = 0;
    for (int i = 0; i < batch.n; ++i) {
      batch.jobs[i] = probs[base + i];
      batch.start[i] = blocks;
      blocks += ceil_div(probs[base + i].M, TS) * ceil_div(probs[base + i].N, TS);
    }
    batch.start[batch.n] = blocks;
    gemm_simt_kernel<<<blocks, NT, 0, stream>>>(batch, dot_acc);
    KFB_CHECK_LAUNCH();
  }
  return KFB_OK;
}

static int run_stage(Context* ctx, std::vector<GemmProb>& probs, float* dot_acc, cudaStream_t stream, int stage = 0) {
  // KFB_CHAIN_MMA=0/1/2 overrides the MMA mode for the precondition chain alone (error attribution experiments)
  const char* chain_env = getenv("KFB_CHAIN_MMA");
  const int chain_mode = chain_env ? atoi(chain_env) : mma_mode();
  // KFB_CHAIN_3X_STAGES="12": in mode 2, only the listed stages (1-4) use 3xTF32 products (experiments)
  const char* only = getenv("KFB_CHAIN_3X_STAGES");
  const bool x3 = chain_mode == 2 && (!only || stage == 0 || strchr(only, '0' + stage) != nullptr);
  if (chain_mode != 0 && ctx->umma_ok) {
    std::vector<GemmProb> tc, simt;
    for (auto& p : probs) (gemm_tc_eligible(p) ? tc : simt).push_back(p);
    int rc = KFB_OK;
    if (!tc.empty()) rc = gemm_grouped_tc(ctx, tc.data(), (int)tc.size(), dot_acc, stream, x3);
    if (rc == KFB_OK && !simt.empty()) rc = gemm_grouped_simt(simt.data(), (int)simt.size(), dot_acc, stream);
    return rc;
  }
  return gemm_grouped_simt(probs.data(), (int)probs.size(), dot_acc, stream);
}

}  // namespace kfb

extern "C" {

int kfb_precondition(const kfb_precond_job* jobs, int32_t n_jobs, float* vg_acc, void* stream_) {
  using namespace kfb;
  Context* ctx;
  int rc = require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(n_jobs >= 0 && (jobs || n_jobs == 0), "bad job array");
  if (n_jobs == 0) return KFB_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  std::vector<GradJob> pack;
  pack.reserve(n_jobs);
  std::vector<GemmProb> s1(n_jobs), s2, s3, s4(n_jobs);
  for (int i = 0; i < n_jobs; ++i) {
    const kfb_precond_job& j = jobs[i];
    KFB_REQUIRE(j.da > 0 && j.dg > 0, "precond job %d: empty", i);
    if (j.use_inverse) {
      KFB_REQUIRE(j.weight_grad && j.EA && j.QG && j.gmat && j.out && j.T, "precond job %d: null pointer", i);
    } else {
      KFB_REQUIRE(j.weight_grad && j.EA && j.QA && j.EG && j.QG && j.dGdA && j.gmat && j.V && j.out && j.T,
                  "precond job %d: null pointer", i);
    }
    KFB_REQUIRE(!j.has_bias || j.bias_grad, "precond job %d: has_bias without bias_grad", i);
    KFB_REQUIRE(j.ld_a >= j.da && j.ld_g >= j.dg && (j.use_inverse || j.ld_dgda >= j.da) && j.ld_v >= j.da &&
                    j.ld_t >= j.dg,
                "precond job %d: pitch too small", i);
    // Without a bias column the contiguous weight gradient already IS the [dg x da] matrix; if
    // its rows are 16-byte aligned the GEMMs read it in place and nothing is packed.
    const float* gm = j.gmat;
    int64_t ld_gm = j.ld_v;
    if (!j.has_bias && (j.da & 3) == 0 && (((uintptr_t)j.weight_grad) & 15) == 0) {
      gm = j.weight_grad;
      ld_gm = j.da;
    } else {
      GradJob g;
      g.wg = (float*)j.weight_grad; g.bg = (float*)j.bias_grad; g.v = j.gmat; g.ld = j.ld_v;
      g.da = j.da; g.dg = j.dg; g.nw = j.da - (j.has_bias ? 1 : 0); g.pad_ = 0;
      pack.push_back(g);
    }
    GemmProb z;
    memset(&z, 0, sizeof(z));
    // T = EA gmat^T          (explicit-inverse mode: EA is A_inv, so T = (gmat A_inv)^T)
    s1[i] = z; s1[i].A = j.EA; s1[i].lda = j.ld_a; s1[i].B = gm; s1[i].ldb = ld_gm;
    s1[i].D = j.T; s1[i].ldd = j.ld_t; s1[i].M = j.da; s1[i].N = j.dg; s1[i].K = j.da;
    if (!j.use_inverse) {
      // V = (EG T^T) * dGdA
      GemmProb p2 = z; p2.A = j.EG; p2.lda = j.ld_g; p2.B = j.T; p2.ldb = j.ld_t;
      p2.D = j.V; p2.ldd = j.ld_v; p2.M = j.dg; p2.N = j.da; p2.K = j.dg;
      p2.Mul = j.dGdA; p2.ldmul = j.ld_dgda;
      s2.push_back(p2);
      // T = QA V^T
      GemmProb p3 = z; p3.A = j.QA; p3.lda = j.ld_a; p3.B = j.V; p3.ldb = j.ld_v;
      p3.D = j.T; p3.ldd = j.ld_t; p3.M = j.da; p3.N = j.dg; p3.K = j.da;
      s3.push_back(p3);
    }
    // out = QG T^T, dot with gmat     (explicit-inverse mode: QG is G_inv)
    s4[i] = z; s4[i].A = j.QG; s4[i].lda = j.ld_g; s4[i].B = j.T; s4[i].ldb = j.ld_t;
    s4[i].D = j.out; s4[i].ldd = j.ld_v; s4[i].M = j.dg; s4[i].N = j.da; s4[i].K = j.dg;
    if (vg_acc) { s4[i].Dot = gm; s4[i].lddot = ld_gm; }
  }
  rc = run_grad<kPack>(pack.data(), (int)pack.size(), nullptr, nullptr, 0.f, 0.f, 1.f, stream);
  if (rc) return rc;
  if ((rc = run_stage(ctx, s1, nullptr, stream, 1))) return rc;
  if (!s2.empty() && (rc = run_stage(ctx, s2, nullptr, stream, 2))) return rc;
  if (!s3.empty() && (rc = run_stage(ctx, s3, nullptr, stream, 3))) return rc;
  return run_stage(ctx, s4, vg_acc, stream, 4);
}

int kfb_gemm_tn(const float* A, int64_t lda, const float* B, int64_t ldb, float* D, int64_t ldd, int32_t M,
                int32_t N, int32_t K, const float* Mul, int64_t ldmul, void* stream) {
  using namespace kfb;
  Context* ctx;
  int rc = require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(A && B && D && M > 0 && N > 0 && K > 0 && lda >= K && ldb >= K && ldd >= N, "bad gemm arguments");
  KFB_REQUIRE(!Mul || ldmul >= N, "bad multiplier pitch");
  std::vector<GemmProb> p(1);
  memset(&p[0], 0, sizeof(GemmProb));
  p[0].A = A; p[0].lda = lda; p[0].B = B; p[0].ldb = ldb; p[0].D = D; p[0].ldd = ldd;
  p[0].M = M; p[0].N = N; p[0].K = K; p[0].Mul = Mul; p[0].ldmul = ldmul;
  return run_stage(ctx, p, nullptr, (cudaStream_t)stream);
}

int kfb_dot_accumulate(const kfb_grad_job* jobs, int32_t n_jobs, float* vg_acc, void* stream) {
  using namespace kfb;
  Context* ctx;
  int rc = require_context(&ctx);
  if (rc) return rc;
  KFB_REQUIRE(vg_acc, "vg_acc is null");
  std::vector<GradJob> g;
  if ((rc = convert_grad_jobs(jobs, n_jobs, &g))) return rc;
  return run_grad<kDot>(g.data(), n_jobs, vg_acc, nullptr, 0.f, 0.f, 1.f, (cudaStream_t)stream);
}

int kfb_scale_writeback(const kfb_grad_job* jobs, int32_t n_jobs, const float* vg_acc, float lr, float kl_clip,
                        float fixed_scale, void* stream) {
  using namespace kfb;
  Context* ctx;
  int rc = require_context(&ctx);
  if (rc) return rc;
  std::vector<GradJob> g;
  if ((rc = convert_grad_jobs(jobs, n_jobs, &g))) return rc;
  return run_grad<kWrite>(g.data(), n_jobs, nullptr, vg_acc, lr, kl_clip, fixed_scale, (cudaStream_t)stream);
}

}  // extern "C"
