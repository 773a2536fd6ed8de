// Grouped TF32 GEMM on the 5th-generation tensor cores:  D[m][n] = sum_k A[m][k] * B[n][k].
//
// One persistent CTA per SM walks the tile list of ALL problems of a stage (all K-FAC layers):
//   warp 0      TMA producer: 128x32 (A) and 128x32 (B) fp32 boxes, 128B-swizzled, 6-stage ring
//   warp 1      MMA issuer: 4 x tcgen05.mma.kind::tf32 (M=128, N=128, K=8) per stage, fp32
//               accumulators in TMEM, double buffered (2 x 128 columns)
//   warps 2-5   epilogue: tcgen05.ld -> registers -> (x dGdA) -> global, optional <D, Dot>
//               warp-shuffle reduction (the KL-clip inner product)
// Operands are consumed as fp32 straight from HBM (kind::tf32 reads the top 19 bits), so there
// is no conversion pass.
#include <algorithm>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "kfb_gemm.cuh"
#include "kfb_umma.cuh"

namespace kfb {
namespace {

constexpr int BM = 128, BN = 128, BK = 32;          // BK floats = 128 bytes = one swizzle atom
constexpr int UMMA_K = 8;
constexpr int kStages = 6;
constexpr int kStageBytes = (BM + BN) * BK * 4;      // 32 KB
constexpr int kThreads = 192;
constexpr int kTmemCols = 2 * BN;                    // double-buffered accumulator
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int kTcBatch = 96;

struct alignas(64) TcProb {
  CUtensorMap mapA, mapB;
  float* D;
  const float* Mul;
  const float* Dot;
  int64_t ldd, ldmul, lddot;
  int M, N, K, tiles_n;
};
struct TcBatch {
  int n;
  int total_tiles;
  int start[kTcBatch + 1];
  TcProb probs[kTcBatch];
};

__device__ __forceinline__ int find_prob(const TcBatch& b, int tile) {
  int lo = 0, hi = b.n;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (b.start[mid] <= tile) lo = mid; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(kThreads, 1) gemm_tc_kernel(const __grid_constant__ TcBatch batch,
                                                            float* __restrict__ dot_acc) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + kStages * kStageBytes);
  uint64_t* empty = full + kStages;
  uint64_t* tfull = empty + kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = (uint32_t*)(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { umma::mbar_init(&full[s], 1); umma::mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { umma::mbar_init(&tfull[s], 1); umma::mbar_init(&tempty[s], 128); }
    umma::fence_barrier_init();
  }
  if (warp == 1) umma::tmem_alloc<kTmemCols>(tmem_slot);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < batch.total_tiles; tile += gridDim.x) {
        const int p = find_prob(batch, tile);
        const TcProb& pr = batch.probs[p];
        const int local = tile - batch.start[p];
        const int m0 = (local / pr.tiles_n) * BM, n0 = (local % pr.tiles_n) * BN;
        const int kblocks = (pr.K + BK - 1) / BK;
        for (int kb = 0; kb < kblocks; ++kb) {
          umma::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + BM * BK * 4;
          umma::mbar_arrive_expect_tx(&full[stage], kStageBytes);
          umma::tma_load_2d(sa, &pr.mapA, &full[stage], kb * BK, m0);
          umma::tma_load_2d(sb, &pr.mapB, &full[stage], kb * BK, n0);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma::make_idesc_tf32(BM, BN);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < batch.total_tiles; tile += gridDim.x) {
        const int p = find_prob(batch, tile);
        const int kblocks = (batch.probs[p].K + BK - 1) / BK;
        umma::mbar_wait(&tempty[acc], acc_phase ^ 1);      // epilogue drained this accumulator
        umma::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          umma::mbar_wait(&full[stage], phase);
          umma::tc_fence_after();
          const uint32_t sa = umma::smem_u32(smem + stage * kStageBytes);
          const uint32_t sb = sa + BM * BK * 4;
          const uint64_t da = umma::make_kmajor_sw128_desc(sa);
          const uint64_t db = umma::make_kmajor_sw128_desc(sb);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advance 8 fp32 = 32 bytes inside the swizzle atom: +2 in the (addr >> 4) field
            umma::mma_tf32_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          }
          umma::tc_commit(&empty[stage]);                  // frees the smem slot when the MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma::tc_commit(&tfull[acc]);                      // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;                                // TMEM lane quarter this warp may read
    int acc = 0; uint32_t acc_phase = 0;
    float dot = 0.f;
    for (int tile = blockIdx.x; tile < batch.total_tiles; tile += gridDim.x) {
      const int p = find_prob(batch, tile);
      const TcProb& pr = batch.probs[p];
      const int local = tile - batch.start[p];
      const int m0 = (local / pr.tiles_n) * BM, n0 = (local % pr.tiles_n) * BN;
      const int m = m0 + q * 32 + lane;
      umma::mbar_wait(&tfull[acc], acc_phase);
      umma::tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        umma::tmem_ld_32x32(taddr + c * 32, r);
        umma::tmem_ld_wait();
        const int nb = n0 + c * 32;
        if (m < pr.M && nb < pr.N) {
          float* drow = pr.D + (int64_t)m * pr.ldd + nb;
          const float* mrow = pr.Mul ? pr.Mul + (int64_t)m * pr.ldmul + nb : nullptr;
          const float* grow = pr.Dot ? pr.Dot + (int64_t)m * pr.lddot + nb : nullptr;
          if (nb + 32 <= pr.N) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                     __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
              if (mrow) {
                float4 mm = *reinterpret_cast<const float4*>(mrow + 4 * j);
                v.x *= mm.x; v.y *= mm.y; v.z *= mm.z; v.w *= mm.w;
              }
              *reinterpret_cast<float4*>(drow + 4 * j) = v;
              if (grow) {
                float4 gg = *reinterpret_cast<const float4*>(grow + 4 * j);
                dot = fmaf(v.x, gg.x, fmaf(v.y, gg.y, fmaf(v.z, gg.z, fmaf(v.w, gg.w, dot))));
              }
            }
          } else {
            const int lim = pr.N - nb;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (j < lim) {
                float v = __uint_as_float(r[j]);
                if (mrow) v *= mrow[j];
                drow[j] = v;
                if (grow) dot = fmaf(v, grow[j], dot);
              }
            }
          }
        }
      }
      umma::tc_fence_before();
      umma::mbar_arrive(&tempty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (dot_acc) {
      dot = warp_sum(dot);
      if (lane == 0 && dot != 0.f) atomicAdd(dot_acc, dot);
    }
  }

  umma::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    umma::tc_fence_after();
    umma::tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// ---- tensor-map cache ----------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct MapKey {
  const void* ptr; int64_t ld; int rows, cols, box_rows;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && ld == o.ld && rows == o.rows && cols == o.cols && box_rows == o.box_rows;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    size_t h = (size_t)k.ptr;
    h = h * 1000003u ^ (size_t)k.ld;
    h = h * 1000003u ^ (size_t)k.rows;
    h = h * 1000003u ^ (size_t)k.cols;
    h = h * 1000003u ^ (size_t)k.box_rows;
    return h;
  }
};
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;
std::mutex g_maps_mu;

// row-major [rows][cols] fp32 matrix, pitch ld floats; box = 32 columns x box_rows rows, SW128
int get_map(Context* ctx, const float* ptr, int64_t ld, int rows, int cols, int box_rows, CUtensorMap* out) {
  MapKey key{ptr, ld, rows, cols, box_rows};
  std::lock_guard<std::mutex> lock(g_maps_mu);
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return KFB_OK; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = ((EncodeTiledFn)ctx->encode_tiled)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptr, dims, strides,
                                                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    KFB_FAIL(KFB_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for %p rows=%d cols=%d ld=%lld", (int)r, ptr, rows, cols,
             (long long)ld);
  if (g_maps.size() > 16384) g_maps.clear();
  g_maps.emplace(key, m);
  *out = m;
  return KFB_OK;
}

}  // namespace

bool gemm_tc_eligible(const GemmProb& p) {
  auto al16 = [](const void* q) { return ((uintptr_t)q & 15) == 0; };
  if (!al16(p.A) || !al16(p.B) || !al16(p.D)) return false;
  if ((p.lda & 3) || (p.ldb & 3) || (p.ldd & 3)) return false;
  if (p.Mul && (!al16(p.Mul) || (p.ldmul & 3))) return false;
  if (p.Dot && (!al16(p.Dot) || (p.lddot & 3))) return false;
  return p.M > 0 && p.N > 0 && p.K > 0;
}

int gemm_grouped_tc(Context* ctx, const GemmProb* probs, int n, float* dot_acc, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    KFB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  static TcBatch batch;   // 32 KB: too large for the stack of some callers; calls are serialised per process
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  for (int base = 0; base < n; base += kTcBatch) {
    batch.n = std::min(kTcBatch, n - base);
    int tiles = 0;
    for (int i = 0; i < batch.n; ++i) {
      const GemmProb& g = probs[base + i];
      TcProb& t = batch.probs[i];
      int rc = get_map(ctx, g.A, g.lda, g.M, g.K, BM, &t.mapA);
      if (rc) return rc;
      rc = get_map(ctx, g.B, g.ldb, g.N, g.K, BN, &t.mapB);
      if (rc) return rc;
      t.D = g.D; t.Mul = g.Mul; t.Dot = g.Dot; t.ldd = g.ldd; t.ldmul = g.ldmul; t.lddot = g.lddot;
      t.M = g.M; t.N = g.N; t.K = g.K; t.tiles_n = ceil_div(g.N, BN);
      batch.start[i] = tiles;
      tiles += ceil_div(g.M, BM) * t.tiles_n;
    }
    batch.start[batch.n] = tiles;
    batch.total_tiles = tiles;
    int grid = std::min(tiles, ctx->num_sms);
    gemm_tc_kernel<<<grid, kThreads, kSmemBytes, stream>>>(batch, dot_acc);
    KFB_CHECK_LAUNCH();
  }
  return KFB_OK;
}

}  // namespace kfb
