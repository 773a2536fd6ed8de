// Grouped TF32 GEMM on the 5th-generation tensor cores:  D[m][n] = sum_k A[m][k] * B[n][k].
//
// Two kernels with the same roles: gemm_tc2_kernel (default) pairs two CTAs per tile with
// tcgen05 cta_group::2 (256 x 256 tiles); gemm_tc_kernel is the single-CTA version (128 x 256).
// One persistent CTA (pair) per SM (pair) walks the tile list of ALL problems of a stage (all layers):
//   warp 0      TMA producer: 128x32 (A) and 128x32 (B) fp32 boxes, 128B-swizzled, 6-stage ring
//   warp 1      MMA issuer: 4 x tcgen05.mma.kind::tf32 (M=128, N=128, K=8) per stage, fp32
//               accumulators in TMEM, double buffered (2 x 128 columns)
//   warps 2-9   epilogue (two warps per TMEM lane quarter, alternating 32-column chunks): tcgen05.ld -> registers -> (x dGdA) -> global, optional <D, Dot>
//               warp-shuffle reduction (the KL-clip inner product)
// Operands are consumed as fp32 straight from HBM (kind::tf32 reads the top 19 bits), so there
// is no conversion pass.
#include <stdlib.h>

#include <algorithm>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "kfb_gemm.cuh"
#include "kfb_umma.cuh"

namespace kfb {
namespace {

// Tile 128 x BN x 32 with BN = 256 when the problem is wide enough, else 128.  fp32 operands are
// heavy on shared-memory bandwidth (TMA writes + tensor-core reads share 128 B/clk): a 128x128
// tile needs 256 B/clk to run the MMA pipe flat out, a 128x256 tile 188 B/clk.
constexpr int BM = 128, BN = 256, BK = 32;          // BK floats = 128 bytes = one swizzle atom
constexpr int UMMA_K = 8;
constexpr int kStages = 4;
constexpr int kStageBytes = (BM + BN) * BK * 4;      // 48 KB
constexpr int kEpilogueWarps = 4;                      // one per TMEM lane quarter (8 spills registers and is slower)
constexpr int kChunkStride = kEpilogueWarps / 4;       // warps sharing a lane quarter interleave chunks
constexpr int kThreads = 32 * (2 + kEpilogueWarps);
constexpr int kTmemCols = 2 * BN;                    // double-buffered accumulator: all 512 columns
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int kTcBatch = 96;

struct alignas(64) TcProb {
  CUtensorMap mapA, mapB;
  float* D;
  const float* Mul;
  const float* Dot;
  int64_t ldd, ldmul, lddot;
  int M, N, K;
  int bn;                     // N extent of this problem's tiles: 128 or 256
  __host__ __device__ int tiles_n() const { return (N + bn - 1) / bn; }
};
struct TcBatch {
  int n;
  int total_tiles;
  int start[kTcBatch + 1];
  TcProb probs[kTcBatch];
};

// Tiles are sorted longest first; handing them out boustrophedon (CTA 0..G-1, then G-1..0, ...)
// instead of round-robin evens out the per-CTA totals without a dynamic scheduler.
__device__ __forceinline__ int snake_tile(int round, int total) {
  const int g = gridDim.x, c = blockIdx.x;
  const int t = round * g + ((round & 1) ? g - 1 - c : c);
  return t < total ? t : -1;      // only the last round can be partial, so -1 also ends the walk
}

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
    for (int s = 0; s < 2; ++s) { umma::mbar_init(&tfull[s], 1); umma::mbar_init(&tempty[s], 32 * kEpilogueWarps); }
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
      for (int round = 0, tile; (tile = snake_tile(round, batch.total_tiles)) >= 0; ++round) {
        const int p = find_prob(batch, tile);
        const TcProb& pr = batch.probs[p];
        const int local = tile - batch.start[p];
        const int m0 = (local / pr.tiles_n()) * BM, n0 = (local % pr.tiles_n()) * pr.bn;
        const int kblocks = (pr.K + BK - 1) / BK;
        const uint32_t tx = (uint32_t)(BM + pr.bn) * BK * 4;
        for (int kb = 0; kb < kblocks; ++kb) {
          umma::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + BM * BK * 4;
          umma::mbar_arrive_expect_tx(&full[stage], tx);
          umma::tma_load_2d(sa, &pr.mapA, &full[stage], kb * BK, m0);
          umma::tma_load_2d(sb, &pr.mapB, &full[stage], kb * BK, n0);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int round = 0, tile; (tile = snake_tile(round, batch.total_tiles)) >= 0; ++round) {
        const int p = find_prob(batch, tile);
        const int kblocks = (batch.probs[p].K + BK - 1) / BK;
        const uint32_t idesc = umma::make_idesc_tf32(BM, batch.probs[p].bn);
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
    for (int round = 0, tile; (tile = snake_tile(round, batch.total_tiles)) >= 0; ++round) {
      const int p = find_prob(batch, tile);
      const TcProb& pr = batch.probs[p];
      const int local = tile - batch.start[p];
      const int m0 = (local / pr.tiles_n()) * BM, n0 = (local % pr.tiles_n()) * pr.bn;
      // Each thread owns one output row (TMEM lane) and walks it 32 columns at a time.  The
      // epilogue operands (dGdA for the Hadamard stage, the gradient for <v,g>) are loaded one
      // chunk AHEAD, the first chunk before the accumulator is even complete, so their HBM
      // latency hides behind the MMAs instead of stalling the TMEM drain (which is what bounds
      // the small-K stages).
      const int m = m0 + q * 32 + lane;
      const bool row_ok = m < pr.M;
      const float* mrow = (pr.Mul && row_ok) ? pr.Mul + (int64_t)m * pr.ldmul + n0 : nullptr;
      const float* grow = (pr.Dot && row_ok) ? pr.Dot + (int64_t)m * pr.lddot + n0 : nullptr;
      float* drow = pr.D + (int64_t)m * pr.ldd + n0;
      const int ncols = min(pr.bn, pr.N - n0);          // valid columns of this tile
      float4 mm[8], gg[8];
      auto load_aux = [&](int c, float4 (&dm)[8], float4 (&dg)[8]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          dm[j] = make_float4(1.f, 1.f, 1.f, 1.f);
          dg[j] = make_float4(0.f, 0.f, 0.f, 0.f);
          const int col = c * 32 + 4 * j;
          if (col + 3 < ncols) {
            if (mrow) dm[j] = __ldg(reinterpret_cast<const float4*>(mrow + col));
            if (grow) dg[j] = __ldg(reinterpret_cast<const float4*>(grow + col));
          } else if (col < ncols) {
            float tm[4] = {1.f, 1.f, 1.f, 1.f}, tg[4] = {0.f, 0.f, 0.f, 0.f};
            for (int e = 0; e < 4; ++e)
              if (col + e < ncols) {
                if (mrow) tm[e] = mrow[col + e];
                if (grow) tg[e] = grow[col + e];
              }
            dm[j] = make_float4(tm[0], tm[1], tm[2], tm[3]);
            dg[j] = make_float4(tg[0], tg[1], tg[2], tg[3]);
          }
        }
      };
      const int half = (warp - 2) >> 2;                  // which of the two warps of this lane quarter
      const int nchunks = (ncols + 31) / 32;
      if (half < nchunks) load_aux(half, mm, gg);
      umma::mbar_wait(&tfull[acc], acc_phase);
      umma::tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = half; c < nchunks; c += kChunkStride) {
        uint32_t r[32];
        umma::tmem_ld_32x32(taddr + c * 32, r);
        float4 mm_next[8], gg_next[8];
        if (c + kChunkStride < nchunks) load_aux(c + kChunkStride, mm_next, gg_next);
        umma::tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int col = c * 32 + 4 * j;
            float4 v = make_float4(__uint_as_float(r[4 * j]) * mm[j].x, __uint_as_float(r[4 * j + 1]) * mm[j].y,
                                   __uint_as_float(r[4 * j + 2]) * mm[j].z, __uint_as_float(r[4 * j + 3]) * mm[j].w);
            if (col + 3 < ncols) {
              *reinterpret_cast<float4*>(drow + col) = v;
            } else if (col < ncols) {
              const float tv[4] = {v.x, v.y, v.z, v.w};
              for (int e = 0; e < 4; ++e)
                if (col + e < ncols) drow[col + e] = tv[e];
            }
            dot = fmaf(v.x, gg[j].x, fmaf(v.y, gg[j].y, fmaf(v.z, gg[j].z, fmaf(v.w, gg[j].w, dot))));
          }
        }
        if (c + kChunkStride < nchunks) {
#pragma unroll
          for (int j = 0; j < 8; ++j) { mm[j] = mm_next[j]; gg[j] = gg_next[j]; }
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

// ==============================================================================================
// CTA-pair variant (cta_group::2): a cluster of two CTAs on one TPC computes a 256 x BN tile.
// Each CTA loads 128 rows of A and HALF of the B tile (BN/2 rows); the pair's tensor cores read
// both halves, so per SM the shared-memory traffic per MMA drops from 12 KB written + 12 KB read
// to 8 KB + 12 KB, lifting the shared-memory ceiling of fp32-operand TF32 from 67 % to ~80 %.
// The leader CTA (cluster rank 0) issues the MMAs; TMA completions of both CTAs are credited to the
// leader's full barrier; commits are multicast to both CTAs' empty / accumulator-full barriers.
// ==============================================================================================
constexpr int kStages2 = 6;
constexpr int kStageBytes2 = (128 + 128) * BK * 4;                 // A 16 KB + half of B 16 KB
constexpr int kSmemBytes2 = kStages2 * kStageBytes2 + 1024 + 256;
// 3xTF32 build of the pair kernel (k3x): a stage also holds the LOW parts of its two tiles, written by four
// extra warps from the fp32 tiles TMA delivered: [A 16 KB][B half 16 KB][A_lo][B_lo], three stages
constexpr int kStages3 = 3;
constexpr int kStageBytes3 = 2 * kStageBytes2;
constexpr int kSplitWarps = 4;

__device__ __forceinline__ int snake_cluster_tile(int round, int total) {
  const int g = gridDim.x >> 1, c = blockIdx.x >> 1;
  const int t = round * g + ((round & 1) ? g - 1 - c : c);
  return t < total ? t : -1;
}

// kEpi: what the epilogue fuses - 0 nothing, 1 the Hadamard multiplier (stage 2), 2 the <D, Dot> reduction
// (stage 4), 3 both (generic).  kEW: epilogue warps, 4 (one per TMEM lane quarter) or 8 (two per quarter,
// alternating 32-column chunks).  Specialising on kEpi halves the registers the operand look-ahead
// holds, which is what lets eight epilogue warps fit without spilling: the short-K stages are bound
// by the drain, not by the MMAs.
// kCoal: the epilogue passes every 32 x 32 chunk through a 4 KB shared-memory buffer of its warp so that its
// global accesses are whole 128-byte row segments (a thread owns an accumulator ROW, so straight from the
// registers one store instruction touches 32 different lines, 16 of every sector's 32 bytes).
// k3x: error-compensated "3xTF32" products.  kind::tf32 reads the top 19 bits of an fp32 operand, i.e. x_hi =
// trunc(x); with x_lo = x - x_hi (exact in fp32) the sum A_lo B_hi + A_hi B_lo + A_hi B_hi, accumulated in fp32,
// carries a relative error of ~2^-20 instead of 2^-10.  The low parts never touch global memory: four extra warps
// compute them from the fp32 tiles in shared memory, stage by stage, between the TMA's arrival and the MMAs
// (fence.proxy.async in between), so the operand traffic out of L2 - the bound of the single-pass kernel - is
// unchanged and only the MMA count triples.  In this build every CTA's TMA reports to its OWN full barrier (its
// split warps wait on it) and the split warps of both CTAs report to the leader's xfull, which the issuer waits on.
template <int kEpi, int kEW, bool kCoal, bool k3x>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(32 * (2 + kEW + (k3x ? kSplitWarps : 0)), 1)
    gemm_tc2_kernel(const __grid_constant__ TcBatch batch, float* __restrict__ dot_acc) {
  constexpr bool kMul = (kEpi & 1) != 0, kDot = (kEpi & 2) != 0;
  constexpr int kStages2 = k3x ? kStages3 : kfb::kStages2;
  constexpr int kStageBytes2 = k3x ? kStageBytes3 : kfb::kStageBytes2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + kStages2 * kStageBytes2);
  uint64_t* empty = full + kStages2;
  uint64_t* tfull = empty + kStages2;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = (uint32_t*)(tempty + 2);
  uint64_t* xfull = (uint64_t*)(tmem_slot + 2);       // [kStages2], k3x only: low parts of both CTAs are in place

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = umma::cluster_ctarank();
  const bool leader = rank == 0;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kStages2; ++s) { umma::mbar_init(&full[s], 1); umma::mbar_init(&empty[s], 1); }
    // tempty lives in the leader: all epilogue threads of the pair arrive there
    for (int s = 0; s < 2; ++s) { umma::mbar_init(&tfull[s], 1); umma::mbar_init(&tempty[s], 2 * 32 * kEW); }
    if (k3x)
      for (int s = 0; s < kStages2; ++s) umma::mbar_init(&xfull[s], 2 * kSplitWarps);
    umma::fence_barrier_init();
  }
  if (warp == 1) umma::tmem_alloc_pair<kTmemCols>(tmem_slot);
  umma::tc_fence_before();
  __syncthreads();
  umma::cluster_sync_all();                 // both CTAs' barriers exist before anything remote happens
  umma::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int round = 0, tile; (tile = snake_cluster_tile(round, batch.total_tiles)) >= 0; ++round) {
        const int p = find_prob(batch, tile);
        const TcProb& pr = batch.probs[p];
        const int local = tile - batch.start[p];
        const int half = pr.bn >> 1;
        const int m0 = (local / pr.tiles_n()) * 256 + (int)rank * 128;
        const int n0 = (local % pr.tiles_n()) * pr.bn + (int)rank * half;
        const int kblocks = (pr.K + BK - 1) / BK;
        const uint32_t tx = 2u * (uint32_t)(128 + half) * BK * 4;      // both CTAs' bytes
        for (int kb = 0; kb < kblocks; ++kb) {
          umma::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes2;
          uint8_t* sb = sa + 128 * BK * 4;
          if (k3x) {
            umma::mbar_arrive_expect_tx(&full[stage], tx >> 1);        // this CTA's own bytes, on its own barrier
            umma::tma_load_2d(sa, &pr.mapA, &full[stage], kb * BK, m0);
            umma::tma_load_2d(sb, &pr.mapB, &full[stage], kb * BK, n0);
          } else {
            if (leader) umma::mbar_arrive_expect_tx(&full[stage], tx);
            umma::tma_load_2d_pair(sa, &pr.mapA, &full[stage], kb * BK, m0);
            umma::tma_load_2d_pair(sb, &pr.mapB, &full[stage], kb * BK, n0);
          }
          if (++stage == kStages2) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader only) =====================
    if (leader && lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
#ifdef KFB_GEMM_PROFILE
      long long pf_tempty = 0, pf_full = 0, pf_issue = 0, pf_tiles = 0, pf_kb = 0;
      const long long pf_t0 = clock64();
#endif
      for (int round = 0, tile; (tile = snake_cluster_tile(round, batch.total_tiles)) >= 0; ++round) {
        const int p = find_prob(batch, tile);
        const int kblocks = (batch.probs[p].K + BK - 1) / BK;
        const uint32_t idesc = umma::make_idesc_tf32(256, batch.probs[p].bn);
#ifdef KFB_GEMM_PROFILE
        const long long pa = clock64();
#endif
        umma::mbar_wait(&tempty[acc], acc_phase ^ 1);
        umma::tc_fence_after();
#ifdef KFB_GEMM_PROFILE
        pf_tempty += clock64() - pa;
        ++pf_tiles;
#endif
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
#ifdef KFB_GEMM_PROFILE
          const long long pb = clock64();
#endif
          if (k3x) umma::mbar_wait_acquire_cluster_a(umma::smem_u32(&xfull[stage]), phase);
          else umma::mbar_wait(&full[stage], phase);
          umma::tc_fence_after();
#ifdef KFB_GEMM_PROFILE
          const long long pc = clock64();
          pf_full += pc - pb;
          ++pf_kb;
#endif
          const uint32_t sa = umma::smem_u32(smem + stage * kStageBytes2);
          const uint32_t sb = sa + 128 * BK * 4;
          const uint64_t da = umma::make_kmajor_sw128_desc(sa);
          const uint64_t db = umma::make_kmajor_sw128_desc(sb);
          if (k3x) {
            // small terms first: A_lo B_hi, A_hi B_lo, then A_hi B_hi (the hardware's truncation IS the hi part)
            const uint64_t dal = umma::make_kmajor_sw128_desc(sa + kfb::kStageBytes2);
            const uint64_t dbl = umma::make_kmajor_sw128_desc(sb + kfb::kStageBytes2);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma::mma_tf32_ss_pair(d_tmem, dal + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma::mma_tf32_ss_pair(d_tmem, da + 2 * k, dbl + 2 * k, idesc, 1u);
          }
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma::mma_tf32_ss_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, k3x ? 1u : (uint32_t)((kb | k) != 0));
          umma::tc_commit_pair(&empty[stage]);
          if (++stage == kStages2) { stage = 0; phase ^= 1; }
#ifdef KFB_GEMM_PROFILE
          pf_issue += clock64() - pc;
#endif
        }
        umma::tc_commit_pair(&tfull[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
#ifdef KFB_GEMM_PROFILE
      if ((blockIdx.x % 36) == 0)
        printf("gemm<%d,%d> cta %3d: total %lld clk, tiles %lld k-blocks %lld | issuer: wait tempty %lld, wait full %lld, issue %lld\n", kEpi, kEW,
               blockIdx.x, clock64() - pf_t0, pf_tiles, pf_kb, pf_tempty, pf_full, pf_issue);
#endif
    }
  } else if (k3x && warp >= 2 + kEW) {
    // ===================== split warps (k3x, both CTAs): x_lo = x - trunc19(x) for this CTA's two tiles =====
    // Elementwise on the swizzled image as it lies (the layout does not matter); warp w takes a quarter of each
    // 16 KB tile.  Rows past the half of a 128-wide B tile hold stale bytes: written, never read by the MMAs.
    const int w = warp - (2 + kEW);
    int stage = 0; uint32_t phase = 0;
    for (int round = 0, tile; (tile = snake_cluster_tile(round, batch.total_tiles)) >= 0; ++round) {
      const int p = find_prob(batch, tile);
      const int kblocks = (batch.probs[p].K + BK - 1) / BK;
      for (int kb = 0; kb < kblocks; ++kb) {
        umma::mbar_wait(&full[stage], phase);                 // this CTA's TMA bytes have landed
        const uint32_t base = umma::smem_u32(smem + stage * kStageBytes2) + (uint32_t)w * 4096u + (uint32_t)lane * 16u;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          uint4 v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i)
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(v[i].x), "=r"(v[i].y), "=r"(v[i].z), "=r"(v[i].w)
                         : "r"(base + (uint32_t)t * 16384u + (uint32_t)i * 512u));
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            // the low part is itself rounded to nearest TF32 here, so the hardware's truncation of it is exact
            auto lo = [](uint32_t x) {
              uint32_t r;
              asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(__uint_as_float(x) - __uint_as_float(x & 0xFFFFE000u)));
              return r;
            };
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(base + (uint32_t)kfb::kStageBytes2 + (uint32_t)t * 16384u + (uint32_t)i * 512u),
                         "r"(lo(v[i].x)), "r"(lo(v[i].y)), "r"(lo(v[i].z)), "r"(lo(v[i].w))
                         : "memory");
          }
        }
        umma::fence_proxy_async_smem();                       // generic-proxy writes -> the MMAs' async-proxy reads
        __syncwarp();
        if (lane == 0) umma::mbar_arrive_leader_release_a(umma::smem_u32(&xfull[stage]));
        if (++stage == kStages2) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (both CTAs: each drains its own 128 accumulator rows) =====
    const int q = warp & 3;                      // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;            // kEW = 8: which of the two warps of this quarter
    constexpr int kStride = kEW / 4;
    int acc = 0; uint32_t acc_phase = 0;
    float dot = 0.f;
#ifdef KFB_GEMM_PROFILE
    long long pe_wait = 0, pe_drain = 0;
#endif
    for (int round = 0, tile; (tile = snake_cluster_tile(round, batch.total_tiles)) >= 0; ++round) {
      const int p = find_prob(batch, tile);
      const TcProb& pr = batch.probs[p];
      const int local = tile - batch.start[p];
      const int m0 = (local / pr.tiles_n()) * 256 + (int)rank * 128, n0 = (local % pr.tiles_n()) * pr.bn;
      const int m = m0 + q * 32 + lane;
      const bool row_ok = m < pr.M;
      const float* xrow = nullptr;               // the fused operand's row: Mul (stage 2) or Dot (stage 4)
      const float* grow = nullptr;               // kEpi == 3 only: Dot beside Mul
      if (kMul && pr.Mul && row_ok) xrow = pr.Mul + (int64_t)m * pr.ldmul + n0;
      if (kDot && pr.Dot && row_ok) (kMul ? grow : xrow) = pr.Dot + (int64_t)m * pr.lddot + n0;
      float* drow = pr.D + (int64_t)m * pr.ldd + n0;
      const int ncols = min(pr.bn, pr.N - n0);
      auto load_row = [&](const float* row, int c, float4 (&d)[8], float fill) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          d[j] = make_float4(fill, fill, fill, fill);
          const int col = c * 32 + 4 * j;
          if (!row) continue;
          if (col + 3 < ncols) {
            d[j] = __ldg(reinterpret_cast<const float4*>(row + col));
          } else if (col < ncols) {
            float t[4] = {fill, fill, fill, fill};
            for (int e = 0; e < 4; ++e)
              if (col + e < ncols) t[e] = row[col + e];
            d[j] = make_float4(t[0], t[1], t[2], t[3]);
          }
        }
      };
      const float xfill = kMul ? 1.f : 0.f;
      const int nchunks = (ncols + 31) / 32;
      if (kCoal) {
        // ---- coalesced drain: lane (rsub = lane >> 3, slot = lane & 7) handles columns 4*slot..4*slot+3 of
        // rows 4*i + rsub, i = 0..7, of the chunk: 8 lanes = one 128-byte row segment
        const uint32_t stg = umma::smem_u32(smem + kStages2 * kStageBytes2 + 256) + (uint32_t)(warp - 2) * 4096u;
        const int rsub = lane >> 3, slot = lane & 7;
        const int mrow0 = m0 + q * 32 + rsub;                    // global row of i = 0
        const float* mul_base = (kMul && pr.Mul) ? pr.Mul + n0 + 4 * slot : nullptr;
        const float* dot_base = (kDot && pr.Dot) ? pr.Dot + n0 + 4 * slot : nullptr;
        float* d_base = pr.D + n0 + 4 * slot;
        auto load_aux = [&](const float* base, int64_t ld, int c, float4 (&d)[8], float fill) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            d[i] = make_float4(fill, fill, fill, fill);
            const int mm = mrow0 + 4 * i, col = c * 32 + 4 * slot;
            if (!base || mm >= pr.M) continue;
            const float* ptr = base + (int64_t)mm * ld + c * 32;
            if (col + 3 < ncols) d[i] = __ldg(reinterpret_cast<const float4*>(ptr));
            else if (col < ncols) {
              float t[4] = {fill, fill, fill, fill};
              for (int e = 0; e < 4; ++e)
                if (col + e < ncols) t[e] = ptr[e];
              d[i] = make_float4(t[0], t[1], t[2], t[3]);
            }
          }
        };
        float4 ax[8], ag[8];
        if (kMul && half < nchunks) load_aux(mul_base, pr.ldmul, half, ax, 1.f);
        if (kDot && half < nchunks) load_aux(dot_base, pr.lddot, half, ag, 0.f);
        umma::mbar_wait(&tfull[acc], acc_phase);
        umma::tc_fence_after();
        const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
        for (int c = half; c < nchunks; c += kStride) {
          uint32_t r[32];
          umma::tmem_ld_32x32(taddr + c * 32, r);
          float4 axn[8], agn[8];
          const bool more = c + kStride < nchunks;
          if (kMul && more) load_aux(mul_base, pr.ldmul, c + kStride, axn, 1.f);
          if (kDot && more) load_aux(dot_base, pr.lddot, c + kStride, agn, 0.f);
          umma::tmem_ld_wait();
          // registers (row = lane) -> shared, 16-byte slots XOR-swizzled by the row: conflict-free both ways
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stg + (uint32_t)lane * 128u + (uint32_t)((j ^ (lane & 7)) << 4)),
                         "r"(r[4 * j]), "r"(r[4 * j + 1]), "r"(r[4 * j + 2]), "r"(r[4 * j + 3])
                         : "memory");
          __syncwarp();
          const int col = c * 32 + 4 * slot;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = 4 * i + rsub, mm = mrow0 + 4 * i;
            float4 v;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                         : "r"(stg + (uint32_t)rr * 128u + (uint32_t)((slot ^ (rr & 7)) << 4)));
            if (kMul) { v.x *= ax[i].x; v.y *= ax[i].y; v.z *= ax[i].z; v.w *= ax[i].w; }
            if (mm < pr.M) {
              float* dp = d_base + (int64_t)mm * pr.ldd + c * 32;
              if (col + 3 < ncols) *reinterpret_cast<float4*>(dp) = v;
              else if (col < ncols) {
                const float tv[4] = {v.x, v.y, v.z, v.w};
                for (int e = 0; e < 4; ++e)
                  if (col + e < ncols) dp[e] = tv[e];
              }
              if (kDot) dot = fmaf(v.x, ag[i].x, fmaf(v.y, ag[i].y, fmaf(v.z, ag[i].z, fmaf(v.w, ag[i].w, dot))));
            }
          }
          __syncwarp();
          if (more) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (kMul) ax[i] = axn[i];
              if (kDot) ag[i] = agn[i];
            }
          }
        }
        umma::tc_fence_before();
        umma::mbar_arrive_leader(&tempty[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        continue;
      }
      float4 xx[8], gg[8];
      if (kEpi != 0 && half < nchunks) load_row(xrow, half, xx, xfill);
      if (kEpi == 3 && half < nchunks) load_row(grow, half, gg, 0.f);
#ifdef KFB_GEMM_PROFILE
      const long long pe0 = clock64();
#endif
      umma::mbar_wait(&tfull[acc], acc_phase);
      umma::tc_fence_after();
#ifdef KFB_GEMM_PROFILE
      const long long pe1 = clock64();
      pe_wait += pe1 - pe0;
#endif
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = half; c < nchunks; c += kStride) {
        uint32_t r[32];
        umma::tmem_ld_32x32(taddr + c * 32, r);
        float4 xn[8], gn[8];
        const bool more = c + kStride < nchunks;
        if (kEpi != 0 && more) load_row(xrow, c + kStride, xn, xfill);
        if (kEpi == 3 && more) load_row(grow, c + kStride, gn, 0.f);
        umma::tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int col = c * 32 + 4 * j;
            float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                   __uint_as_float(r[4 * j + 3]));
            if (kMul) { v.x *= xx[j].x; v.y *= xx[j].y; v.z *= xx[j].z; v.w *= xx[j].w; }
            if (col + 3 < ncols) {
              *reinterpret_cast<float4*>(drow + col) = v;
            } else if (col < ncols) {
              const float tv[4] = {v.x, v.y, v.z, v.w};
              for (int e = 0; e < 4; ++e)
                if (col + e < ncols) drow[col + e] = tv[e];
            }
            if (kEpi == 2) dot = fmaf(v.x, xx[j].x, fmaf(v.y, xx[j].y, fmaf(v.z, xx[j].z, fmaf(v.w, xx[j].w, dot))));
            if (kEpi == 3) dot = fmaf(v.x, gg[j].x, fmaf(v.y, gg[j].y, fmaf(v.z, gg[j].z, fmaf(v.w, gg[j].w, dot))));
          }
        }
        if (more) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (kEpi != 0) xx[j] = xn[j];
            if (kEpi == 3) gg[j] = gn[j];
          }
        }
      }
      umma::tc_fence_before();
#ifdef KFB_GEMM_PROFILE
      pe_drain += clock64() - pe1;
#endif
      umma::mbar_arrive_leader(&tempty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
#ifdef KFB_GEMM_PROFILE
    if (warp == 2 && lane == 0 && (blockIdx.x % 36) == 0)
      printf("gemm<%d,%d> cta %3d epilogue warp: wait tfull %lld, drain %lld\n", kEpi, kEW, blockIdx.x, pe_wait, pe_drain);
#endif
    if (kDot && dot_acc) {
      dot = warp_sum(dot);
      if (lane == 0 && dot != 0.f) atomicAdd(dot_acc, dot);
    }
  }

  umma::tc_fence_before();
  __syncthreads();
  umma::cluster_sync_all();                 // the peer may still be signalling this CTA's barriers
  if (warp == 1) {
    umma::tc_fence_after();
    umma::tmem_dealloc_pair<kTmemCols>(tmem_base);
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

typedef void (*Tc2Kernel)(const TcBatch, float*);
template <int kEpi, int kEW, bool kCoal, bool k3x = false>
static int launch_tc2(Context* ctx, const TcBatch& batch, int clusters, float* dot_acc, cudaStream_t stream) {
  static bool attr_set[64] = {};        // the opt-in to > 48 KB of dynamic shared memory is per device (and per instantiation)
  // + one 32 x 32 fp32 staging buffer per epilogue warp
  constexpr int smem = (k3x ? kStages3 * kStageBytes3 + 1024 + 256 : kSmemBytes2) + (kCoal ? kEW * 4096 : 0);
  static_assert(smem <= 232448, "shared memory of the GEMM kernel exceeds 227 KB");
  if (!attr_set[ctx->device]) {
    KFB_CUDA(cudaFuncSetAttribute(gemm_tc2_kernel<kEpi, kEW, kCoal, k3x>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set[ctx->device] = true;
  }
  gemm_tc2_kernel<kEpi, kEW, kCoal, k3x><<<2 * clusters, 32 * (2 + kEW + (k3x ? kSplitWarps : 0)), smem, stream>>>(batch, dot_acc);
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

static int gemm_grouped_tc2(Context* ctx, const GemmProb* probs, int n, float* dot_acc, cudaStream_t stream, bool x3) {
  // eight epilogue warps by default (KFB_GEMM_EPI_WARPS=4: one per TMEM lane quarter)
  static const int epi_warps = []() { const char* e = getenv("KFB_GEMM_EPI_WARPS"); return e && atoi(e) == 4 ? 4 : 8; }();
  std::vector<GemmProb> sorted(probs, probs + n);
  std::stable_sort(sorted.begin(), sorted.end(), [](const GemmProb& a, const GemmProb& b) { return a.K > b.K; });
  probs = sorted.data();
  // what the epilogue of this launch has to fuse (a stage's problems all ask for the same thing)
  int epi = 0;
  for (int i = 0; i < n; ++i) epi |= (probs[i].Mul ? 1 : 0) | ((probs[i].Dot && dot_acc) ? 2 : 0);
  static TcBatch batch;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  for (int base = 0; base < n; base += kTcBatch) {
    batch.n = std::min(kTcBatch, n - base);
    int tiles = 0;
    for (int i = 0; i < batch.n; ++i) {
      const GemmProb& g = probs[base + i];
      TcProb& t = batch.probs[i];
      int rc = get_map(ctx, g.A, g.lda, g.M, g.K, 128, &t.mapA);
      if (rc) return rc;
      t.bn = g.N > 128 ? 256 : 128;
      rc = get_map(ctx, g.B, g.ldb, g.N, g.K, t.bn / 2, &t.mapB);      // each CTA loads half of the B tile
      if (rc) return rc;
      t.D = g.D; t.Mul = g.Mul; t.Dot = dot_acc ? g.Dot : nullptr; t.ldd = g.ldd; t.ldmul = g.ldmul; t.lddot = g.lddot;
      t.M = g.M; t.N = g.N; t.K = g.K;
      batch.start[i] = tiles;
      tiles += ceil_div(g.M, 256) * t.tiles_n();
    }
    batch.start[batch.n] = tiles;
    batch.total_tiles = tiles;
    const int clusters = std::min(tiles, ctx->num_sms / 2);
    static const bool coal = []() { const char* e = getenv("KFB_GEMM_COALESCE"); return !e || atoi(e) != 0; }();
    int rc;
    if (x3) {            // error-compensated products: always the coalesced eight-warp build
      rc = epi == 0 ? launch_tc2<0, 8, true, true>(ctx, batch, clusters, dot_acc, stream)
           : epi == 1 ? launch_tc2<1, 8, true, true>(ctx, batch, clusters, dot_acc, stream)
           : epi == 2 ? launch_tc2<2, 8, true, true>(ctx, batch, clusters, dot_acc, stream)
                      : launch_tc2<3, 8, true, true>(ctx, batch, clusters, dot_acc, stream);
    } else if (coal) {   // coalesced drain through shared memory, eight warps
      rc = epi == 0 ? launch_tc2<0, 8, true>(ctx, batch, clusters, dot_acc, stream)
           : epi == 1 ? launch_tc2<1, 8, true>(ctx, batch, clusters, dot_acc, stream)
           : epi == 2 ? launch_tc2<2, 8, true>(ctx, batch, clusters, dot_acc, stream)
                      : launch_tc2<3, 8, true>(ctx, batch, clusters, dot_acc, stream);
    } else if (epi_warps == 8) {
      rc = epi == 0 ? launch_tc2<0, 8, false>(ctx, batch, clusters, dot_acc, stream)
           : epi == 1 ? launch_tc2<1, 8, false>(ctx, batch, clusters, dot_acc, stream)
           : epi == 2 ? launch_tc2<2, 8, false>(ctx, batch, clusters, dot_acc, stream)
                      : launch_tc2<3, 8, false>(ctx, batch, clusters, dot_acc, stream);
    } else {
      rc = epi == 0 ? launch_tc2<0, 4, false>(ctx, batch, clusters, dot_acc, stream)
           : epi == 1 ? launch_tc2<1, 4, false>(ctx, batch, clusters, dot_acc, stream)
           : epi == 2 ? launch_tc2<2, 4, false>(ctx, batch, clusters, dot_acc, stream)
                      : launch_tc2<3, 4, false>(ctx, batch, clusters, dot_acc, stream);
    }
    if (rc) return rc;
  }
  return KFB_OK;
}

int gemm_grouped_tc(Context* ctx, const GemmProb* probs, int n, float* dot_acc, cudaStream_t stream, bool x3) {
  // CTA-pair kernel by default (measured 0.72 -> 0.67 ms for the ResNet-50 chain; routing the
  // problems with M <= 128 to the single-CTA kernel instead was slower); KFB_GEMM_2CTA=0 selects
  // the single-CTA kernel
  static const bool use_pair = []() { const char* e = getenv("KFB_GEMM_2CTA"); return !e || atoi(e) != 0; }();
  if (use_pair || x3) return gemm_grouped_tc2(ctx, probs, n, dot_acc, stream, x3);
  // longest tiles first: with the static round-robin tile schedule this is close to LPT
  std::vector<GemmProb> sorted(probs, probs + n);
  std::stable_sort(sorted.begin(), sorted.end(), [](const GemmProb& a, const GemmProb& b) { return a.K > b.K; });
  probs = sorted.data();
  static TcBatch batch;   // 32 KB: too large for the stack of some callers; calls are serialised per process
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  static bool attr_set[64] = {};
  if (!attr_set[ctx->device]) {
    KFB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set[ctx->device] = true;
  }
  for (int base = 0; base < n; base += kTcBatch) {
    batch.n = std::min(kTcBatch, n - base);
    int tiles = 0;
    for (int i = 0; i < batch.n; ++i) {
      const GemmProb& g = probs[base + i];
      TcProb& t = batch.probs[i];
      int rc = get_map(ctx, g.A, g.lda, g.M, g.K, BM, &t.mapA);
      if (rc) return rc;
      t.bn = g.N > 128 ? 256 : 128;
      rc = get_map(ctx, g.B, g.ldb, g.N, g.K, t.bn, &t.mapB);
      if (rc) return rc;
      t.D = g.D; t.Mul = g.Mul; t.Dot = g.Dot; t.ldd = g.ldd; t.ldmul = g.ldmul; t.lddot = g.lddot;
      t.M = g.M; t.N = g.N; t.K = g.K;
      batch.start[i] = tiles;
      tiles += ceil_div(g.M, BM) * t.tiles_n();
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
