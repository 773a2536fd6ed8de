// Factor SYRK on the 5th-generation tensor cores:  F += w * X^T X  for one 256x256 block (2x2
// tiles of 128) of the upper triangle and one K chunk per work item.  X is the im2col-unfolded
// activation (kfac/layers/conv.py:50-70) or the NCHW output gradient viewed as (N*S) x C; the
// unfolded matrix is never materialised.  Blocking 2x2 halves the operand traffic per FLOP (4 tiles
// feed 4 tile products instead of 2 tiles per product) and the shared-memory traffic per MMA.
//
//   warps 0-3   epilogue: tcgen05.ld -> x weight -> red.global.add.v4.f32 into the job's scratch
//               (laid out the way the accumulators are held, DevJob::native), or the fused finish
//   warp  0     lane 0 is also the MMA issuer: tcgen05.mma.kind::tf32, M=128, N=256 (or 128), K=8;
//               the block's four fp32 accumulators fill all 512 TMEM columns, so issuing and
//               draining alternate anyway; a diagonal block loads its two tiles once for both
//               operands and skips the lower-left product.  (20 warps = 5 per scheduler leave 96
//               registers per thread; a 21st warp would cap every thread at 80.)
//   warps 4-19  producers: bring X^T tiles (128 features x 32 rows of X, K-major, 128-byte swizzle)
//               into shared memory, never through registers unless the element type forces it:
//               - fp32 NCHW operands with 16-byte image rows and >= 32 pixels per image: one lane
//                 issues a TMA box (32 pixels x 128 channels x 1 image) per tile and stage; K runs
//                 over per-image 32-pixel blocks, the tail of an image's last block is zero-filled
//                 by the TMA unit.  NCHW operands TMA cannot read (7x7 maps, 16-bit captures) are
//                 first copied to fp32 [channels][images x pixels] (kfb_factor.cu, PadJob variant 4)
//               - kxk convs: from the per-kx stride-compacted zero-padded staging planes (PadJob):
//                 one 16-byte cp.async.cg per lane (four output pixels of one tap), a whole stage per
//                 lane in flight, completion reported by cp.async.mbarrier.arrive - no border masks,
//                 no waiting for data
//               - fallback, capture in place (KFB_SYRK_STAGE=0 / shapes the staging refuses): 4-byte
//                 cp.async with zero fill under a per-lane validity bitmask over (ky,kx); 16-bit
//                 captures in place are widened in registers
// Operands stay fp32 in shared memory; kind::tf32 reads the top 19 bits, so no conversion pass.
#include <algorithm>
#include <cstdlib>
#include <type_traits>
#include <vector>

#include "kfb_factor.cuh"
#include "kfb_umma.cuh"

namespace kfb {
namespace {

constexpr int TM = 128;                 // tile side (features)
constexpr int BK = 32;                  // rows of X per stage = one 128-byte swizzle row
constexpr int UMMA_K = 8;
constexpr int kTileBytes = TM * BK * 4; // 16 KB
constexpr int kProducerWarps = 16;
constexpr int kRowsPerWarp = TM / kProducerWarps;   // 8 rows of each tile per producer warp
constexpr int kTmemCols = 512;
constexpr int kSched = 4;                 // ring of item indices handed from the fetching lane to the CTA's warps

// Two builds of the same kernel:
//   single CTA : a stage holds the block's four tiles (64 KB), 3 stages; the four accumulators fill
//                TMEM, so MMA issue (warp 0, lane 0) and the drain (warps 0-3) alternate; 20 warps.
//   CTA pair   : cta_group::2 - the leader's MMAs use both SMs; each CTA gathers ONE row tile and ONE
//                column tile (32 KB per stage, 6 stages: twice the bytes in flight per FLOP) and
//                holds 128 x 256 accumulators, double-buffered, so the drain of one item overlaps the
//                MMAs of the next; warp 4 issues (leader) or relays "my half has landed" (peer).
template <bool kPair>
struct Cfg {
  static constexpr int kStages = kPair ? 6 : 3;
  static constexpr int kSlots = kPair ? 2 : 4;                 // 16 KB tiles per stage
  static constexpr int kStageBytes = kSlots * kTileBytes;
  static constexpr int kFirstProducer = kPair ? 5 : 4;         // first producer warp
  static constexpr int kThreads = 32 * (kFirstProducer + kProducerWarps);
  static constexpr int kBarOff = kStages * kStageBytes;
  static constexpr int kSmemBytes = kBarOff + 1024 + 256;       // 1 KB alignment slack + barriers
};

// A tile row (= one feature of X) is described by ONE packed word: (element offset << 6) | bit,
// where `bit` selects the bit of the lane's validity mask that says whether this feature exists
// at the lane's X row.  Bits 0..48 are the (ky,kx) positions of an im2col window (image border /
// padding), kBitAlways is set whenever the X row exists, kBitNever is never set (features >= d,
// and the bias feature, which is written separately).
constexpr uint32_t kBitAlways = 62;
constexpr uint32_t kBitNever = 63;

__device__ __forceinline__ uint32_t row_info(const DevSegment& s, int d, int i) {
  if (i >= d || (s.has_bias && i == d - 1)) return kBitNever;
  if (s.kind == KFB_SEG_NCHW) return ((uint32_t)(i * s.h * s.w) << 6) | kBitAlways;
  int khw = s.kh * s.kw;
  int ci = i / khw;
  int q = i - ci * khw;
  int ky = q / s.kw, kx = q - ky * s.kw;
  // staging copy (PadJob): one h x w plane per (channel, kx), rows are input rows
  if (s.nomask) return ((uint32_t)(((ci * s.kw + kx) * s.h + ky) * s.w) << 6) | kBitAlways;
  return ((uint32_t)((ci * s.h + ky) * s.w + kx) << 6) | (uint32_t)q;
}

// Does the feature described by `info` exist for the lane whose validity mask is (mlo, mhi)?
// kWide = false: all bits the segment can use live in one 32-bit word (im2col windows of up to
// 25 taps keep their bits in mlo; kBitAlways / kBitNever are bits 30 / 31 of mhi remapped by the
// caller), so the test is one wrapping funnel shift.
template <bool kWide>
__device__ __forceinline__ bool row_valid(uint32_t info, uint32_t mlo, uint32_t mhi) {
  if (kWide) {
    uint32_t m = (info & 32u) ? mhi : mlo;
    return (__funnelshift_r(m, 0u, info) & 1u) != 0;       // shift amount taken mod 32
  }
  return (__funnelshift_r(mlo, 0u, info) & 1u) != 0;
}

// x / d by the segment's precomputed multiplier (see magic_div in kfb_factor.cu); x < 2^31
__device__ __forceinline__ uint32_t fast_div(uint32_t x, uint32_t mul, uint32_t shift) {
  return mul ? (__umulhi(x, mul) >> shift) : x;
}

// Items are sorted longest first; hand them out boustrophedon (CTA 0..G-1, then G-1..0, ...) so the
// per-CTA totals even out without a dynamic scheduler.
template <bool kPair>
__device__ __forceinline__ int snake_item(int round, int total) {
  const int g = kPair ? gridDim.x >> 1 : gridDim.x, c = kPair ? blockIdx.x >> 1 : blockIdx.x;
  const int t = round * g + ((round & 1) ? g - 1 - c : c);
  return t < total ? t : -1;
}
// The single-CTA build walks a CONTIGUOUS range of the item list instead (cta_first[blockIdx.x] ..
// cta_first[blockIdx.x + 1]): the host cut the linearised work into one equal share per CTA, so a
// CTA's consecutive K pieces of one block are one item and the accumulators stay in TMEM across them.
template <bool kPair>
__device__ __forceinline__ int next_item(int round, int total, const int* __restrict__ cta_first) {
  if (kPair || cta_first == nullptr) return snake_item<kPair>(round, total);
  const int it = cta_first[blockIdx.x] + round;
  return it < cta_first[blockIdx.x + 1] ? it : -1;
}

// explicit shared-space stores: through a generic pointer the compiler emits ST.E (generic
// address translation on every store) instead of STS
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void sts_v4(uint32_t addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Asynchronous global -> shared copies (LDGSTS): no register staging, so a lane keeps a whole stage
// in flight instead of waiting for every batch of loads before it may store.  src_bytes = 0 writes
// zeros without touching global memory (borders, padding rows, K tails).
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4_zfill(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
// this thread's arrival on `bar` happens when all of its earlier cp.async copies have landed
__device__ __forceinline__ void cp_async_arrive(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

// 256-bit global accesses (sm_100: LDG.256 / STG.256), 32-byte aligned
__device__ __forceinline__ void ldg_v8(const float* p, float4& a, float4& b) {
  asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
               : "l"(p)
               : "memory");
}
__device__ __forceinline__ void stg_v8(float* p, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1,
                                       uint32_t b2, uint32_t b3) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0),
               "r"(b1), "r"(b2), "r"(b3)
               : "memory");
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// ---- last-arriver finish of one 32 x 32 chunk of a split-K job (DevJob::la) --------------------------------
// The chunk's complete sum is read back from the scratch in the accumulator's own layout (lane = row i, 32
// columns per lane; coalesced), the scratch is re-zeroed for the next update, and
//   new = (old * alpha/(1-alpha) + rescale*sum) * (1-alpha)           (kfac/layers/utils.py:175-178)
// goes to state[i][j] and state[j][i].  `old` is read from the MIRROR position state[j][i] (the state is kept
// exactly symmetric), which the 32 lanes of a warp touch as one 128-byte segment, as do the mirror stores; the
// row stores are one full line per lane.  Diagonal tiles hold j >= i only.
__device__ __forceinline__ void la_finish_chunk(const DevJob& job, int tile_m, int tile_n, int q, int c, int lane, float ratio,
                                                float one_m, float resc, uint8_t* stage) {
  const int d = job.d;
  const int jb = tile_n * TM + c * 32;
  const bool diag_tile = tile_m == tile_n;
  if (jb >= d || tile_m * TM + q * 32 >= d || (diag_tile && c < q)) return;       // warp-uniform: nothing was accumulated here
  const int i = tile_m * TM + q * 32 + lane;
  float* cb = job.scratch + (((int64_t)(tile_m * job.T128 + tile_n) * 16 + q * 4 + c) << 10) + (lane << 2);
  uint32_t r[32];
#pragma unroll
  for (int g = 0; g < 8; ++g)
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[4 * g]), "=r"(r[4 * g + 1]), "=r"(r[4 * g + 2]), "=r"(r[4 * g + 3]) : "l"(cb + g * 128) : "memory");
#pragma unroll
  for (int g = 0; g < 8; ++g) __stcg(reinterpret_cast<float4*>(cb + g * 128), make_float4(0.f, 0.f, 0.f, 0.f));
  const bool row_ok = i < d;
  const bool keep_old = resc < 0.f;                       // every capture was dropped: the average stays
  const bool on_diag = diag_tile && c == q;               // the chunk the diagonal runs through: j >= i only
  float* mcol = job.state + (int64_t)jb * job.ld_state + i;                      // state[jb + e][i]
  const bool full = !on_diag && jb + 31 < d && row_ok;
  // two halves of 16 columns: all of a half's old values are requested before the first is used (the stores
  // below go through the same pointer, so the compiler would otherwise wait for each load in turn)
#pragma unroll
  for (int h = 0; h < 32; h += 16) {
    float o[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      const int j = jb + h + e;
      const bool ok = full || (row_ok && j < d && (!on_diag || j >= i));
      o[e] = 0.f;
      if (job.seed_identity) o[e] = i == j ? 1.f : 0.f;
      else if (ok) o[e] = __ldcg(mcol + (int64_t)(h + e) * job.ld_state);
    }
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      const int j = jb + h + e;
      const bool ok = full || (row_ok && j < d && (!on_diag || j >= i));
      const float v = keep_old ? o[e] : (o[e] * ratio + __uint_as_float(r[h + e]) * resc) * one_m;
      r[h + e] = __float_as_uint(v);
      if (ok && job.write_state && (!diag_tile || j > i)) __stcg(mcol + (int64_t)(h + e) * job.ld_state, v);
    }
  }
  // Row stores.  A lane owns a row, so a store instruction of the warp touches 32 different lines with 16 bytes
  // each - measured at ~13 bytes per clock and SM.  With a 4 KB staging buffer (XOR-swizzled 16-byte groups,
  // conflict-free both ways) the chunk is turned so that eight lanes write one whole 128-byte row segment.
  const bool whole = !on_diag && jb + 31 < d;                 // warp-uniform
  if (stage && whole && job.write_state && (job.ld_state & 3) == 0 && (((uintptr_t)job.state) & 15) == 0) {
#pragma unroll
    for (int g = 0; g < 8; ++g)
      *reinterpret_cast<uint4*>(stage + lane * 128 + ((g ^ (lane & 7)) << 4)) = make_uint4(r[4 * g], r[4 * g + 1], r[4 * g + 2], r[4 * g + 3]);
    __syncwarp();
    const int i0 = tile_m * TM + q * 32, g = lane & 7;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int row = 4 * k + (lane >> 3);
      const uint4 v = *reinterpret_cast<const uint4*>(stage + row * 128 + ((g ^ (row & 7)) << 4));
      if (i0 + row < d) __stcg(reinterpret_cast<uint4*>(job.state + (int64_t)(i0 + row) * job.ld_state + jb + 4 * g), v);
    }
    __syncwarp();
    if (!row_ok) return;
  } else if (!row_ok) {
    return;
  } else if (job.write_state) {
    float* srow = job.state + (int64_t)i * job.ld_state + jb;
    if (full && (((uintptr_t)srow) & 15) == 0) {
#pragma unroll
      for (int g = 0; g < 8; ++g)
        *reinterpret_cast<uint4*>(srow + 4 * g) = make_uint4(r[4 * g], r[4 * g + 1], r[4 * g + 2], r[4 * g + 3]);
    } else {
#pragma unroll
      for (int e = 0; e < 32; ++e)
        if (jb + e < d && (!on_diag || jb + e >= i)) srow[e] = __uint_as_float(r[e]);
    }
  }
  if (job.tri_out) {
    float* trow = job.tri_out + ((int64_t)i * d - ((int64_t)i * (i - 1)) / 2 - i);
#pragma unroll
    for (int e = 0; e < 32; ++e)
      if (jb + e < d && jb + e >= i) trow[jb + e] = __uint_as_float(r[e]);
  }
}

// All chunks of one completed 256 x 256 block, shared among `n_warps` warps (this one is number `w`).
__device__ __forceinline__ void la_finish_block(const DevJob& job, const TcItem& item, int w, int n_warps, int lane, float alpha,
                                                uint8_t* stage) {
  const float ratio = alpha / (1.f - alpha), one_m = 1.f - alpha;
  const float resc = job.dev_rescale ? __ldg(job.dev_rescale) : 1.f;
  const int units = item.mi * item.nj * 16;
  for (int u = w; u < units; u += n_warps) {
    const int t = u >> 4, m = t / item.nj, nn = t - m * item.nj;
    if (item.a0 == item.b0 && nn < m) continue;
    la_finish_chunk(job, item.a0 + m, item.b0 + nn, (u >> 2) & 3, u & 3, lane, ratio, one_m, resc, stage);
  }
}

template <bool kPair>
__global__ void __launch_bounds__(Cfg<kPair>::kThreads, 1) syrk_tc_kernel(const DevJob* __restrict__ jobs,
                                                                        const DevSegment* __restrict__ segs,
                                                                        const TcItem* __restrict__ items, int n_items, float alpha,
                                                                        const CUtensorMap* __restrict__ tmaps,
                                                                        const int* __restrict__ cta_first, int* la_ctl, int dyn) {
  using C = Cfg<kPair>;
  constexpr int kStages = C::kStages, kStageBytes = C::kStageBytes, kSlots = C::kSlots;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + C::kBarOff);
  uint64_t* empty = full + kStages;
  uint64_t* tfull = empty + kStages;       // [2]; the single-CTA build uses [0] only
  uint64_t* tempty = tfull + 2;            // [2]; in a pair the LEADER's copy collects both CTAs' arrivals
  uint32_t* tmem_slot = (uint32_t*)(tempty + 2);
  int* la_bcast = (int*)(tmem_slot + 2);              // end of kernel: the completed block all warps finish next
  uint64_t* sched_full = (uint64_t*)(smem + C::kBarOff + 128);   // [kSched] dynamic item hand-out, see fetch_item
  uint64_t* sched_empty = sched_full + kSched;
  volatile int* sched_item = (volatile int*)(sched_empty + kSched);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = kPair ? umma::cluster_ctarank() : 0u;
  const bool leader = rank == 0;

  if (threadIdx.x == 0) {
    // full: every producer thread arrives (cp.async completion or plain); the pair's leader also
    // takes one arrival per stage from the peer's relay
    for (int s = 0; s < kStages; ++s) {
      umma::mbar_init(&full[s], kProducerWarps * 32 + ((kPair && leader) ? 1 : 0));
      umma::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) { umma::mbar_init(&tfull[s], 1); umma::mbar_init(&tempty[s], kPair ? 256 : 128); }
    for (int s = 0; s < kSched; ++s) { umma::mbar_init(&sched_full[s], 1); umma::mbar_init(&sched_empty[s], C::kThreads / 32); }
    umma::fence_barrier_init();
  }
  if (kPair) { if (warp == 4) umma::tmem_alloc_pair<kTmemCols>(tmem_slot); }
  else       { if (warp == 0) umma::tmem_alloc<kTmemCols>(tmem_slot); }
  umma::tc_fence_before();
  __syncthreads();
  if (kPair) umma::cluster_sync_all();      // both CTAs' barriers exist before anything remote happens
  umma::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // ---- which item next?  Default: the static boustrophedon deal / stream-K shares of next_item.  With dyn != 0
  // (single-CTA build, control block present; la_ctl[3] = items handed out so far) every CTA takes the next
  // unclaimed item of the longest-first list when it is ready for one: one lane claims, a small ring of
  // mbarriers hands the index to all warps of the CTA (each is one consumer).
  const bool dynamic = !kPair && la_ctl != nullptr && cta_first == nullptr && dyn != 0;
  auto fetch_item = [&](int round) -> int {
    if (kPair || !dynamic) return next_item<kPair>(round, n_items, cta_first);
    const int slot = round & (kSched - 1);
    const uint32_t ph = (uint32_t)(round / kSched) & 1u;
    if (warp == C::kFirstProducer && lane == 0) {
      umma::mbar_wait(&sched_empty[slot], ph ^ 1u);                        // all warps have read the previous tenant
      int it = round == 0 ? (int)blockIdx.x : (int)gridDim.x + atomicAdd(&la_ctl[3], 1);
      if (it >= n_items) it = -1;
      sched_item[slot] = it;
      umma::mbar_arrive(&sched_full[slot]);
    }
    umma::mbar_wait(&sched_full[slot], ph);
    const int it = sched_item[slot];
    __syncwarp();
    if (lane == 0) umma::mbar_arrive(&sched_empty[slot]);
    return it;
  };

  const uint32_t full_a = umma::smem_u32(full), empty_a = umma::smem_u32(empty);
  if (warp >= C::kFirstProducer) {
    // ===================== producers =====================
    const int pw = warp - C::kFirstProducer;
    // byte offset of this lane inside any of this warp's tile rows: rows are pw + 16*j, so the
    // swizzle term (row & 7) is the same for all of them
    const uint32_t lane_off = (uint32_t)((((lane >> 2) ^ (pw & 7)) << 4) | ((lane & 3) << 2));
    const uint32_t vec_kq = lane & 7, vec_rsub = lane >> 3;
    const uint32_t smem_base = umma::smem_u32(smem);
    int stage = 0; uint32_t phase = 0;
    for (int round = 0, it; (it = fetch_item(round)) >= 0; ++round) {
      const TcItem item = items[it];
      const DevSegment s = segs[item.seg];
      const int d = jobs[item.job].d;
      const bool diag = item.a0 == item.b0;
      // smem slots to fill.  Single CTA: row tiles a0.. (slots 0..mi-1) and, off the diagonal, column
      // tiles b0.. (slots 2..2+nj-1).  Pair: this CTA's row tile a0+rank (slot 0) and, off the
      // diagonal, its half of the column tiles, b0+rank (slot 1); tiles past the factor gather zeros.
      int n_slots;
      int slot_of[kSlots], tile_of[kSlots];
      if (kPair) {
        n_slots = diag ? 1 : 2;
        slot_of[0] = 0; tile_of[0] = item.a0 + (int)rank;
        slot_of[1] = 1; tile_of[1] = item.b0 + (int)rank;
      } else {
        n_slots = diag ? item.mi : item.mi + item.nj;
#pragma unroll
        for (int t = 0; t < kSlots; ++t) {
          const bool is_a = t < item.mi;
          slot_of[t] = is_a ? t : 2 + (t - item.mi);
          tile_of[t] = is_a ? item.a0 + t : item.b0 + (t - item.mi);
        }
      }
      const int S = s.oh * s.ow;
      const uint32_t uS = (uint32_t)S, uow = (uint32_t)s.ow;
      const uint32_t k_end32 = item.k_end;       // K < 2^31 (checked on the host)
      if (s.tma >= 0) {
        // ---- TMA-fed operand: k counts rows of the per-image padded matrix (nb blocks of 32)
        const CUtensorMap* tm = tmaps + s.tma;
        const bool issuer = pw == 0 && lane == 0;
        // the descriptor was written by a host copy into memory that held other descriptors before
        if (issuer) umma::fence_tensormap_acquire(tm);
        const uint32_t unb = (uint32_t)s.nb;
        uint32_t b = item.k_begin / BK;
        uint32_t n = b / unb, sb = b - n * unb;
        for (uint32_t k0 = item.k_begin; k0 < k_end32; k0 += BK) {
          umma::mbar_wait_a(empty_a + 8 * stage, phase ^ 1);
          if (issuer) {
            const uint32_t st = smem_base + stage * kStageBytes;
            if (kPair) {
              // both CTAs' boxes are credited to the LEADER's barrier, which expects the pair's bytes
              if (leader) umma::mbar_arrive_expect_tx_a(full_a + 8 * stage, (uint32_t)(2 * n_slots * kTileBytes));
              else umma::mbar_arrive_a(full_a + 8 * stage);
#pragma unroll
              for (int t = 0; t < kSlots; ++t)
                if (t < n_slots)
                  umma::tma_load_3d_pair_a(st + slot_of[t] * kTileBytes, tm, full_a + 8 * stage, (int)(sb * BK), tile_of[t] * TM, (int)n);
            } else {
              umma::mbar_arrive_expect_tx_a(full_a + 8 * stage, (uint32_t)(n_slots * kTileBytes));
#pragma unroll
              for (int t = 0; t < kSlots; ++t)
                if (t < n_slots)
                  umma::tma_load_3d_a(st + slot_of[t] * kTileBytes, tm, full_a + 8 * stage, (int)(sb * BK), tile_of[t] * TM, (int)n);
            }
          } else {
            umma::mbar_arrive_a(full_a + 8 * stage);
          }
          __syncwarp();
          if (++sb == unb) { sb = 0; ++n; }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        continue;
      }
      const bool vec4 = s.kind == KFB_SEG_NCHW && (S & 3) == 0 &&
                        (((uintptr_t)s.data) & (s.dtype == KFB_F32 ? 15u : 7u)) == 0;
      const bool wide = s.kind == KFB_SEG_IM2COL && s.kh * s.kw > 30;
      // does this warp own the bias feature's row in one of the tiles?
      int bias_slot = -1, bias_row = 0;
      if (s.has_bias) {
#pragma unroll
        for (int t = 0; t < kSlots; ++t) {
          int local = (d - 1) - tile_of[t] * TM;
          if (t < n_slots && local >= 0 && local < TM && (local % kProducerWarps) == pw) {
            bias_slot = slot_of[t];
            bias_row = local;
          }
        }
      }

      // The gather loops are instantiated once per element width so that the fp32 path carries no
      // run-time dtype tests (they cost ~30 % of the factor stage when left in the loop).
      auto gather = [&](auto esh_tag) {
        constexpr int esh = decltype(esh_tag)::value;     // log2(bytes per captured element)
        if (vec4) {
          // ---- NCHW, 16B-alignable: 8 lanes x float4 = one tile row (32 k), 4 rows per instruction
          uint32_t info[kSlots][2];
  #pragma unroll
          for (int t = 0; t < kSlots; ++t)
  #pragma unroll
            for (int i = 0; i < 2; ++i)
              info[t][i] = t < n_slots ? row_info(s, d, tile_of[t] * TM + pw + kProducerWarps * (4 * i + vec_rsub)) : kBitNever;
          for (uint32_t k0 = item.k_begin; k0 < k_end32; k0 += BK) {
            umma::mbar_wait_a(empty_a + 8 * stage, phase ^ 1);
            const uint32_t st = smem_base + stage * kStageBytes;
            const uint32_t k = k0 + 4 * vec_kq;
            const bool kin = k < k_end32;
            const uint32_t n = fast_div(k, s.mul_s, s.sh_s);
            const char* src = reinterpret_cast<const char*>(s.data) + (((int64_t)n * s.c * S + (k - n * uS)) << esh);
            float4 v[kSlots][2];
  #pragma unroll
            for (int t = 0; t < kSlots; ++t)
  #pragma unroll
              for (int i = 0; i < 2; ++i) {
                v[t][i] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (kin && (info[t][i] & 63u) == kBitAlways) {
                  const char* p = src + ((uint64_t)(info[t][i] >> 6) << esh);
                  if (esh == 2) {
                    v[t][i] = __ldg(reinterpret_cast<const float4*>(p));
                  } else {               // four 16-bit captures in one 8-byte load
                    const uint2 h = __ldg(reinterpret_cast<const uint2*>(p));
                    v[t][i] = make_float4(widen16((unsigned short)(h.x & 0xffff), s.dtype), widen16((unsigned short)(h.x >> 16), s.dtype),
                                          widen16((unsigned short)(h.y & 0xffff), s.dtype), widen16((unsigned short)(h.y >> 16), s.dtype));
                  }
                }
              }
  #pragma unroll
            for (int t = 0; t < kSlots; ++t)
  #pragma unroll
              for (int i = 0; i < 2; ++i)
                if (t < n_slots) {
                  const int r = pw + kProducerWarps * (4 * i + vec_rsub);
                  sts_v4(st + slot_of[t] * kTileBytes + r * 128 + ((vec_kq ^ (r & 7)) << 4), v[t][i]);
                }
            umma::fence_proxy_async_smem();
            umma::mbar_arrive_a(full_a + 8 * stage);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        } else if (esh == 2 && s.nomask) {
          // ---- im2col from the staging copy (PadJob): for a tap (ky,kx) four consecutive output pixels
          // of a row are 16 contiguous, aligned bytes, and every tap of every (padded) pixel is in range.
          // 8 lanes x 16 bytes = one tile row (32 k), 4 rows per instruction, like the NCHW path; the
          // copies are asynchronous, so a lane has the whole stage in flight.
          // Word = byte offset of the row's (channel, kx) plane + ky rows | 1 if the feature does not
          // exist (>= d, or the bias feature, written separately below).
          uint32_t off[kSlots][2];
          uint32_t clean = 0xFu;            // bit t: every row of slot t is an existing, non-bias feature
  #pragma unroll
          for (int t = 0; t < kSlots; ++t) {
  #pragma unroll
            for (int i = 0; i < 2; ++i) {
              const uint32_t inf = t < n_slots ? row_info(s, d, tile_of[t] * TM + pw + kProducerWarps * (4 * i + vec_rsub)) : kBitNever;
              off[t][i] = (inf & 63u) == kBitNever ? 1u : ((inf >> 6) << 2);
            }
            if (t < n_slots && (tile_of[t] + 1) * TM > d - s.has_bias) clean &= ~(1u << t);
          }
          const int64_t img = (int64_t)s.c * s.kw * s.h * s.w;
          for (uint32_t k0 = item.k_begin; k0 < k_end32; k0 += BK) {
            umma::mbar_wait_a(empty_a + 8 * stage, phase ^ 1);
            const uint32_t st = smem_base + stage * kStageBytes;
            const uint32_t k = k0 + 4 * vec_kq;           // this lane's four padded pixels: k .. k+3
            const bool kin = k < k_end32;                 // item ends are multiples of 4
            const bool full_stage = k0 + BK <= k_end32;
            const uint32_t n = fast_div(k, s.mul_s, s.sh_s);
            const uint32_t r = k - n * uS;
            const uint32_t oy = fast_div(r, s.mul_w, s.sh_w), ox = r - oy * uow;
            // lanes past the end of the item read image 0 (in range) and are zero-filled below
            const char* src = reinterpret_cast<const char*>(s.data) +
                              ((kin ? (int64_t)n * img : 0) + (int64_t)(oy * s.sh) * s.w + ox) * 4;
  #pragma unroll
            for (int t = 0; t < kSlots; ++t) {
              if (t >= n_slots) break;
              const uint32_t dst = st + slot_of[t] * kTileBytes;
  #pragma unroll
              for (int i = 0; i < 2; ++i) {
                const int rr = pw + kProducerWarps * (4 * i + vec_rsub);
                const uint32_t da = dst + rr * 128 + ((vec_kq ^ (rr & 7)) << 4);
                const uint32_t o = off[t][i];
                if (slot_of[t] == bias_slot && rr == bias_row) continue;     // written below, not zero-filled
                if (full_stage && ((clean >> t) & 1u)) cp_async16(da, src + o);
                else cp_async16_zfill(da, src + (o & ~3u), (kin && !(o & 1u)) ? 16u : 0u);
              }
            }
            // the bias feature is the constant 1 wherever the X row exists (not on the pitch padding)
            if (bias_slot >= 0) {
              const uint32_t kb = k0 + lane;
              const uint32_t nb_ = fast_div(kb, s.mul_s, s.sh_s);
              const uint32_t rb = kb - nb_ * uS;
              const uint32_t xb = rb - fast_div(rb, s.mul_w, s.sh_w) * uow;
              float one = (kb < k_end32 && xb < (uint32_t)s.ow_valid) ? 1.f : 0.f;
              if (s.fmt) {
                // 16-bit planes: a word is two pixels of the row; ow_valid counts PIXELS here
                const uint32_t h1 = s.fmt == KFB_F16 ? 0x3C00u : 0x3F80u;
                const uint32_t p0 = 2u * xb;
                one = __uint_as_float(kb < k_end32 ? ((p0 < (uint32_t)s.ow_valid ? h1 : 0u) | (p0 + 1u < (uint32_t)s.ow_valid ? h1 << 16 : 0u)) : 0u);
              }
              sts_f32(st + bias_slot * kTileBytes + bias_row * 128 + (uint32_t)((((lane >> 2) ^ (bias_row & 7)) << 4) | ((lane & 3) << 2)), one);
              umma::fence_proxy_async_smem();
            }
            cp_async_arrive(full_a + 8 * stage);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        } else {
          // ---- general gather: one X row per lane, this warp's 8 rows of every tile per stage
          uint32_t info[kSlots][kRowsPerWarp];
  #pragma unroll
          for (int t = 0; t < kSlots; ++t)
  #pragma unroll
            for (int j = 0; j < kRowsPerWarp; ++j)
              info[t][j] = t < n_slots ? row_info(s, d, tile_of[t] * TM + pw + kProducerWarps * j) : kBitNever;
          for (uint32_t k0 = item.k_begin; k0 < k_end32; k0 += BK) {
            umma::mbar_wait_a(empty_a + 8 * stage, phase ^ 1);
            const uint32_t st = smem_base + stage * kStageBytes + pw * 128 + lane_off;
            const uint32_t k = k0 + lane;
            const bool kin = k < k_end32;
            const uint32_t n = fast_div(k, s.mul_s, s.sh_s);
            const uint32_t r = k - n * uS;
            const char* src;
            uint32_t mlo = 0, mhi = 0;
            if (s.kind == KFB_SEG_NCHW) {
              src = reinterpret_cast<const char*>(s.data) + (((int64_t)n * s.c * S + r) << esh);
              if (kin) mhi = 1u << (kBitAlways - 32);
            } else {
              const int oy = (int)fast_div(r, s.mul_w, s.sh_w), ox = (int)(r - (uint32_t)oy * uow);
              const int y0 = oy * s.sh - s.ph, x0 = ox * s.sw - s.pw;
              src = reinterpret_cast<const char*>(s.data) + (((int64_t)n * s.c * s.h * s.w + (int64_t)y0 * s.w + x0) << esh);
              if (kin) {
                const int kx_lo = max(0, -x0), kx_hi = min(s.kw, s.w - x0);
                const int ky_lo = max(0, -y0), ky_hi = min(s.kh, s.h - y0);
                uint64_t m = 0;
                const uint64_t xm = kx_hi > kx_lo ? (((1ull << (kx_hi - kx_lo)) - 1) << kx_lo) : 0;
                for (int ky = ky_lo; ky < ky_hi; ++ky) m |= xm << (ky * s.kw);
                mlo = (uint32_t)m;
                mhi = (uint32_t)(m >> 32);
              }
            }
            // narrow: (ky,kx) bits in mlo; "always" moves to bit 30 so one word decides everything
            // (row_info's 6-bit field: 62 & 31 = 30, 63 & 31 = 31 which is never set)
            const uint32_t mnarrow = mlo | ((mhi >> (kBitAlways - 32)) << 30);
            if (esh == 2) {
              // fp32: asynchronous copies, zero-filled where the mask says the tap is outside the image
  #pragma unroll
              for (int t = 0; t < kSlots; ++t)
  #pragma unroll
                for (int j = 0; j < kRowsPerWarp; ++j)
                  if (t < n_slots && !(slot_of[t] == bias_slot && pw + kProducerWarps * j == bias_row)) {
                    const uint32_t inf = info[t][j];
                    const bool ok = wide ? row_valid<true>(inf, mlo, mhi) : row_valid<false>(inf, mnarrow, 0u);
                    cp_async4_zfill(st + slot_of[t] * kTileBytes + j * (kProducerWarps * 128), src + ((uint64_t)(inf >> 6) << 2),
                                    ok ? 4u : 0u);
                  }
            } else {
              // 16-bit captures are widened in registers: two tiles (16 loads per lane) in flight at a time
  #pragma unroll
              for (int tp = 0; tp < kSlots / 2; ++tp) {
                if (2 * tp >= n_slots) break;
                float v[2][kRowsPerWarp];
  #pragma unroll
                for (int u = 0; u < 2; ++u)
  #pragma unroll
                  for (int j = 0; j < kRowsPerWarp; ++j) {
                    const uint32_t inf = info[2 * tp + u][j];
                    const bool ok = wide ? row_valid<true>(inf, mlo, mhi) : row_valid<false>(inf, mnarrow, 0u);
                    v[u][j] = 0.f;
                    if (ok) v[u][j] = widen16(__ldg(reinterpret_cast<const unsigned short*>(src + ((uint64_t)(inf >> 6) << 1))), s.dtype);
                  }
  #pragma unroll
                for (int u = 0; u < 2; ++u)
  #pragma unroll
                  for (int j = 0; j < kRowsPerWarp; ++j)
                    if (2 * tp + u < n_slots)
                      sts_f32(st + slot_of[2 * tp + u] * kTileBytes + j * (kProducerWarps * 128), v[u][j]);
              }
            }
            // the bias feature is the constant 1 wherever the X row exists
            if (bias_slot >= 0)
              sts_f32(smem_base + stage * kStageBytes + bias_slot * kTileBytes + bias_row * 128 + lane_off,
                      kin ? 1.f : 0.f);
            if (esh != 2 || bias_slot >= 0) umma::fence_proxy_async_smem();
            cp_async_arrive(full_a + 8 * stage);       // with no copies in flight this is an immediate arrival
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      };
      if (s.dtype == KFB_F32) gather(std::integral_constant<int, 2>{});
      else gather(std::integral_constant<int, 1>{});
    }
  } else if (kPair && warp == 4) {
    // ===================== pair: MMA issuer (leader) / landing relay (peer) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int round = 0, it; (it = fetch_item(round)) >= 0; ++round) {
        const TcItem item = items[it];
        if (leader) {
          const bool diag = item.a0 == item.b0;
          umma::mbar_wait(&tempty[acc], acc_phase ^ 1);      // both CTAs have drained this accumulator
          umma::tc_fence_after();
          // rows: tile a0 (this CTA) and a0+1 (peer); columns: tile b0 (here) and b0+1 (peer's half)
          const uint32_t idesc = umma::make_idesc_tf32(2 * TM, 2 * TM);
          const uint32_t d_tmem = tmem_base + acc * 256;
          bool first = true;
          for (uint32_t k0 = item.k_begin; k0 < item.k_end; k0 += BK) {
            umma::mbar_wait_a(full_a + 8 * stage, phase);
            umma::tc_fence_after();
            const uint32_t sbase = umma::smem_u32(smem + stage * kStageBytes);
            const uint64_t da = umma::make_kmajor_sw128_desc(sbase);
            const uint64_t db = umma::make_kmajor_sw128_desc(sbase + (diag ? 0 : kTileBytes));
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma::mma_tf32_ss_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, (first && k == 0) ? 0u : 1u);
            first = false;
            umma::tc_commit_pair(&empty[stage]);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
          umma::tc_commit_pair(&tfull[acc]);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        } else {
          for (uint32_t k0 = item.k_begin; k0 < item.k_end; k0 += BK) {
            // This CTA's copies of the stage have landed in its shared memory (the local barrier says
            // so); tell the leader with a plain remote arrive.  No cluster-scope release/acquire pair:
            // nobody READS the data through the generic proxy on the other side - the pair's tensor
            // cores read it in place - and ptxas turns release.cluster into MEMBAR.GPU + ERRBAR per
            // stage (and acquire.cluster into an L1 invalidate), which made the relay the bottleneck.
            umma::mbar_wait_a(full_a + 8 * stage, phase);
            umma::mbar_arrive_leader_a(full_a + 8 * stage);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else {
    // ============ epilogue (warps 0..3); single-CTA build: warp 0, lane 0 also issues the MMAs ============
    const int q = warp;
    uint32_t acc_phase = 0;
    int acc = 0;
    int stage = 0; uint32_t phase = 0;
#ifdef KFB_SYRK_PROFILE
    // clock64 breakdown of the issuing thread's time (build this file with -DKFB_SYRK_PROFILE)
    long long pf_tempty = 0, pf_full = 0, pf_issue = 0, pf_drain = 0, pf_items = 0, pf_stages = 0, pf_arrive = 0;
    long long pf_type_clk[3] = {0, 0, 0}, pf_type_units[3] = {0, 0, 0};
    const long long pf_t0 = clock64();
#endif
    for (int round = 0, it; (it = fetch_item(round)) >= 0; ++round) {
      const TcItem item = items[it];
      const DevJob& job = jobs[item.job];
#ifdef KFB_SYRK_PROFILE
      const long long pf_item0 = clock64();
#endif
      // device-side multiplier of the segment weight (AMP unscale; 0 = capture dropped, kfb_capture_filter)
      const float* dscale = segs[item.seg].dev_scale;
      const float dmul = dscale ? __ldg(dscale) : 1.f;
      // a fused job has no finish pass to apply its rescale (< 0: no capture survived, keep the state)
      const float fresc = (job.fused && job.dev_rescale) ? __ldg(job.dev_rescale) : 1.f;
      const bool dropped = dmul == 0.f || fresc < 0.f;    // warp-uniform: skip every write of this item
      const float weight = segs[item.seg].weight * dmul * fresc;
      const int d = job.d;
      const bool diag = item.a0 == item.b0;
      const float ratio = alpha / (1.f - alpha), one_m = 1.f - alpha;
      // ---- finish of one 32-row x 32-column chunk whose COMPLETE sum is in r[] (thread = row i):
      //   new = (old * alpha/(1-alpha) + w*acc) * (1-alpha)     (kfac/layers/utils.py:175-178)
      // written to the state row, mirrored into column i (one coalesced 128-byte store per column: the 32
      // lanes are 32 consecutive rows), and exported to the packed triangle.  Used by the fused jobs (one
      // item holds the whole sum; a diagonal TILE then holds both of its triangles) and by the last-arriver
      // finish of split-K jobs (upper_only on diagonal tiles: the scratch holds j >= i).
      auto finish_chunk = [&](int i, int jb, uint32_t (&r)[32], bool diag_tile, bool upper_only, float weight, bool keep_old) {
        const bool row_ok = i < d;
        if (row_ok) {
          const float* orow = job.state + (int64_t)i * job.ld_state + jb;
          if (!job.seed_identity && jb + 31 < d && (((uintptr_t)orow) & 15) == 0) {
            // this thread's 32 old values are one 128-byte line: eight 16-byte loads (a scalar
            // load per value costs 32 cache wavefronts per instruction, one per lane's row)
            float4 o[8];
            if ((((uintptr_t)orow) & 31) == 0) {
#pragma unroll
              for (int g = 0; g < 8; g += 2) ldg_v8(orow + 4 * g, o[g], o[g + 1]);
            } else {
#pragma unroll
              for (int g = 0; g < 8; ++g) o[g] = *reinterpret_cast<const float4*>(orow + 4 * g);
            }
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              r[4 * g] = __float_as_uint(keep_old ? o[g].x : (o[g].x * ratio + __uint_as_float(r[4 * g]) * weight) * one_m);
              r[4 * g + 1] = __float_as_uint(keep_old ? o[g].y : (o[g].y * ratio + __uint_as_float(r[4 * g + 1]) * weight) * one_m);
              r[4 * g + 2] = __float_as_uint(keep_old ? o[g].z : (o[g].z * ratio + __uint_as_float(r[4 * g + 2]) * weight) * one_m);
              r[4 * g + 3] = __float_as_uint(keep_old ? o[g].w : (o[g].w * ratio + __uint_as_float(r[4 * g + 3]) * weight) * one_m);
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const int j = jb + e;
              float old = 0.f;
              if (j < d) old = job.seed_identity ? (i == j ? 1.f : 0.f) : orow[e];
              r[e] = __float_as_uint(keep_old ? old : (old * ratio + __uint_as_float(r[e]) * weight) * one_m);
            }
          }
          if (job.write_state && upper_only) {
            // the block's sum was accumulated for j >= i only (split-K scratch): write that part of the row;
            // the rest of it arrives through the mirror stores of the rows above
            float* srow = job.state + (int64_t)i * job.ld_state + jb;
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (jb + e < d && jb + e >= i) srow[e] = __uint_as_float(r[e]);
          } else if (job.write_state) {
            float* srow = job.state + (int64_t)i * job.ld_state + jb;
            if (jb + 31 < d && (((uintptr_t)srow) & 31) == 0) {
#pragma unroll
              for (int g = 0; g < 8; g += 2)
                stg_v8(srow + 4 * g, r[4 * g], r[4 * g + 1], r[4 * g + 2], r[4 * g + 3], r[4 * g + 4], r[4 * g + 5],
                       r[4 * g + 6], r[4 * g + 7]);
            } else
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const int j = jb + 4 * g;
              if (j + 3 < d) {
                *reinterpret_cast<uint4*>(srow + 4 * g) = make_uint4(r[4 * g], r[4 * g + 1], r[4 * g + 2], r[4 * g + 3]);
              } else {
                for (int e = 0; e < 4; ++e)
                  if (j + e < d) srow[4 * g + e] = __uint_as_float(r[4 * g + e]);
              }
            }
          }
          if (job.tri_out) {
            float* trow = job.tri_out + ((int64_t)i * d - ((int64_t)i * (i - 1)) / 2 - i);
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const int j = jb + e;
              if (j < d && j >= i) trow[j] = __uint_as_float(r[e]);
            }
          }
        }
        if ((!diag_tile || upper_only) && job.write_state) {
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const int j = jb + e;
            if (row_ok && j < d && (!upper_only || j > i)) job.state[(int64_t)j * job.ld_state + i] = __uint_as_float(r[e]);
          }
        }
      };
      if (!kPair && warp == 0) {
        if (lane == 0) {
#ifdef KFB_SYRK_PROFILE
          const long long pa = clock64();
#endif
          umma::mbar_wait(tempty, acc_phase ^ 1);            // all four warps have drained TMEM
          umma::tc_fence_after();
#ifdef KFB_SYRK_PROFILE
          pf_tempty += clock64() - pa;
          ++pf_items;
#endif
          // Per item: everything about the (up to two) MMAs of a k step except the stage offset.  Row
          // tile m runs against column tiles n0..nj-1 in ONE instruction (N = 128 or 256); on the
          // diagonal the column tiles are the row tiles and products below the diagonal are skipped.
          uint32_t idesc_m[2], a_lo[2], b_lo[2], d_tm[2];
          const int fmt16 = segs[item.seg].fmt;      // 16-bit TMA-fed operand: kind::f16 over the same tile bytes
          const uint32_t lo0 = umma::kmajor_sw128_desc_lo(umma::smem_u32(smem));
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            const int n0 = diag ? m : 0;
            idesc_m[m] = fmt16 ? umma::make_idesc_f16(fmt16 == KFB_BF16 ? 1 : 0, TM, (item.nj - n0) * TM)
                               : umma::make_idesc_tf32(TM, (item.nj - n0) * TM);
            a_lo[m] = lo0 + ((m * kTileBytes) >> 4);
            b_lo[m] = lo0 + ((((diag ? 0 : 2) + n0) * kTileBytes) >> 4);
            d_tm[m] = tmem_base + m * 256 + n0 * TM;
          }
          const bool two = item.mi > 1;
          uint32_t acc_flag = 0;
          for (uint32_t k0 = item.k_begin; k0 < item.k_end; k0 += BK) {
#ifdef KFB_SYRK_PROFILE
            const long long pb = clock64();
#endif
            umma::mbar_wait_a(full_a + 8 * stage, phase);
            umma::tc_fence_after();
#ifdef KFB_SYRK_PROFILE
            const long long pc = clock64();
            pf_full += pc - pb;
            ++pf_stages;
#endif
            const uint32_t soff = (uint32_t)stage * (kStageBytes >> 4);
            if (fmt16) {
#pragma unroll
              for (int k = 0; k < BK / UMMA_K; ++k)
                umma::mma_f16_ss_lohi(d_tm[0], a_lo[0] + soff + 2 * k, b_lo[0] + soff + 2 * k, idesc_m[0], k == 0 ? acc_flag : 1u);
              if (two) {
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k)
                  umma::mma_f16_ss_lohi(d_tm[1], a_lo[1] + soff + 2 * k, b_lo[1] + soff + 2 * k, idesc_m[1], k == 0 ? acc_flag : 1u);
              }
            } else {
#pragma unroll
              for (int k = 0; k < BK / UMMA_K; ++k)
                umma::mma_tf32_ss_lohi(d_tm[0], a_lo[0] + soff + 2 * k, b_lo[0] + soff + 2 * k, idesc_m[0], k == 0 ? acc_flag : 1u);
              if (two) {
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k)
                  umma::mma_tf32_ss_lohi(d_tm[1], a_lo[1] + soff + 2 * k, b_lo[1] + soff + 2 * k, idesc_m[1], k == 0 ? acc_flag : 1u);
              }
            }
            acc_flag = 1u;
            umma::tc_commit(&empty[stage]);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
#ifdef KFB_SYRK_PROFILE
            pf_issue += clock64() - pc;
#endif
          }
          umma::tc_commit(tfull);
        }
        __syncwarp();
      }
      umma::mbar_wait(&tfull[acc], acc_phase);
      umma::tc_fence_after();
#ifdef KFB_SYRK_PROFILE
      const long long pe = clock64();
#endif
      // single CTA: both row tiles, accumulators at columns m*256; pair: this CTA's row tile only
      // (a0 + rank), accumulator buffer `acc` at columns acc*256
      const int m_begin = kPair ? (int)rank : 0;
      const int m_end = kPair ? min((int)rank + 1, item.mi) : item.mi;
      for (int m = m_begin; m < m_end; ++m) {
        const int i = (item.a0 + m) * TM + q * 32 + lane;
        for (int nn = diag ? m : 0; nn < item.nj; ++nn) {
          const int j0 = (item.b0 + nn) * TM;
          const uint32_t taddr = tmem_base + (kPair ? acc : m) * 256 + nn * TM + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
          for (int c = 0; c < TM / 32; ++c) {
            const int jb = j0 + c * 32;
            if (jb >= d) break;                               // warp-uniform
            if (dropped) break;
            uint32_t r[32];
            umma::tmem_ld_32x32(taddr + c * 32, r);
            umma::tmem_ld_wait();
            if (job.fused) {
              finish_chunk(i, jb, r, (item.a0 + m) == (item.b0 + nn), false, weight, false);
              continue;
            }
            if (i < d && jb + 31 >= i) {
              // scratch in the accumulator's own layout (DevJob::native): the 32 lanes of one red.v4
              // are 512 contiguous bytes
              float* cb = job.scratch + (((int64_t)((item.a0 + m) * job.T128 + (item.b0 + nn)) * 16 + q * 4 + c) << 10) + (lane << 2);
#pragma unroll
              for (int g = 0; g < 8; ++g) {
                const int j = jb + 4 * g;
                if (j + 3 < d && j >= i) {
                  red_add_v4(cb + g * 128, __uint_as_float(r[4 * g]) * weight, __uint_as_float(r[4 * g + 1]) * weight,
                             __uint_as_float(r[4 * g + 2]) * weight, __uint_as_float(r[4 * g + 3]) * weight);
                } else {
#pragma unroll
                  for (int e = 0; e < 4; ++e)
                    if (j + e < d && j + e >= i) atomicAdd(cb + g * 128 + e, __uint_as_float(r[4 * g + e]) * weight);
                }
              }
            }
          }
        }
      }
#ifdef KFB_SYRK_PROFILE
      const long long pla = clock64();
#endif
      if (!kPair && job.la) {
        // ---- last-arriver finish (split-K jobs): every item of a 256 x 256 block counts itself in after its
        // reds; the one that completes the block publishes it in the queue of completed blocks.  CTAs that have
        // run out of items take blocks from that queue and finish them with all their warps (end of the kernel)
        // - off every MMA issuer's path, and in the time the early CTAs would otherwise idle.
        __threadfence();                                     // this thread's reds before the count
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 0) {
          int* cnt = job.arrive + (item.a0 >> 1) * ((job.T128 + 1) >> 1) + (item.b0 >> 1);
          if (atomicAdd(cnt, 1) == job.la_items - 1) {
            *cnt = 0;                                        // ready for the next update
            __threadfence();
            const int slot = atomicAdd(&la_ctl[0], 1);
            atomicExch(&la_ctl[kLaCtlInts + slot], it + 1);
          }
        }
      }
#ifdef KFB_SYRK_PROFILE
      pf_arrive += clock64() - pla;
#endif
      umma::tc_fence_before();
#ifdef KFB_SYRK_PROFILE
      {
        const long long now = clock64();
        pf_drain += now - pe;
        const DevSegment& ps = segs[item.seg];
        const int ty = ps.tma >= 0 ? 0 : ps.nomask ? 1 : 2;
        const long long units = item.a0 == item.b0 ? (item.mi * (item.mi + 1)) / 2 : item.mi * item.nj;
        pf_type_clk[ty] += now - pf_item0;
        pf_type_units[ty] += units * ((item.k_end - item.k_begin) / BK);
      }
#endif
      if (kPair) {
        umma::mbar_arrive_leader(&tempty[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      } else {
        umma::mbar_arrive(tempty);
        acc_phase ^= 1;
      }
    }
#ifdef KFB_SYRK_PROFILE
    if (!kPair && threadIdx.x == 0 && (blockIdx.x % 21) == 0)
      printf("syrk cta %3d: total %lld clk, items %lld stages %lld | issuer: wait tempty %lld, wait full %lld, issue %lld | drain %lld (arrival %lld) | "
             "clk per stage-unit: tma %.0f (%lld), staged %.0f (%lld), gather %.0f (%lld)\n",
             blockIdx.x, clock64() - pf_t0, pf_items, pf_stages, pf_tempty, pf_full, pf_issue, pf_drain, pf_arrive,
             pf_type_units[0] ? (double)pf_type_clk[0] / pf_type_units[0] : 0.0, pf_type_units[0],
             pf_type_units[1] ? (double)pf_type_clk[1] / pf_type_units[1] : 0.0, pf_type_units[1],
             pf_type_units[2] ? (double)pf_type_clk[2] / pf_type_units[2] : 0.0, pf_type_units[2]);
#endif
  }

  umma::tc_fence_before();
  __syncthreads();
  if (!kPair && la_ctl != nullptr) {
    // Finish completed blocks (running average + mirror) for as long as the queue holds any: never waits for a
    // block that is not complete yet - whoever completes it later drains the queue after its own items, so the
    // last CTA to get here leaves it empty.  The stage ring is free by now: 4 KB of it per warp turn the chunks.
#ifdef KFB_SYRK_PROFILE
    const long long pf0 = clock64();
    int pf_n = 0;
#endif
    for (;;) {
      if (threadIdx.x == 0) {
        int got = -1;
        for (;;) {
          const int popped = *(volatile int*)&la_ctl[1], pushed = *(volatile int*)&la_ctl[0];
          if (popped >= pushed) break;
          if (atomicCAS(&la_ctl[1], popped, popped + 1) != popped) continue;
          volatile int* slot = la_ctl + kLaCtlInts + popped;
          while ((got = *slot) == 0) __nanosleep(64);        // its publisher is between two instructions
          *slot = 0;                                         // empty for the next update
          got -= 1;
          __threadfence();
          break;
        }
        *la_bcast = got;
      }
      __syncthreads();
      const int got = *la_bcast;
      __syncthreads();
      if (got < 0) break;
      const TcItem item = items[got];
      la_finish_block(jobs[item.job], item, warp, C::kThreads / 32, lane, alpha, smem + warp * 4096);
#ifdef KFB_SYRK_PROFILE
      ++pf_n;
#endif
    }
    // the last CTA out resets the queue's counters (every push and pop of this launch has happened by then)
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(&la_ctl[2], 1) == (int)gridDim.x - 1) { la_ctl[0] = 0; la_ctl[1] = 0; la_ctl[2] = 0; la_ctl[3] = 0; }
    }
#ifdef KFB_SYRK_PROFILE
    if (threadIdx.x == 0 && (blockIdx.x % 21) == 0) printf("syrk cta %3d: finished %d blocks in %lld clk\n", blockIdx.x, pf_n, clock64() - pf0);
#endif
  }
  if (kPair) {
    umma::cluster_sync_all();                 // the peer may still be signalling this CTA's barriers
    if (warp == 4) {
      umma::tc_fence_after();
      umma::tmem_dealloc_pair<kTmemCols>(tmem_base);
    }
  } else if (warp == 0) {
    umma::tc_fence_after();
    umma::tmem_dealloc<kTmemCols>(tmem_base);
  }
}

}  // namespace

// Build the tensor-core work lists for the jobs marked use_tc.
//
// Default: (2x2 tile block, K chunk of up to 4096 rows) items, longest first, handed out boustrophedon.
// KFB_SYRK_STREAMK=1 (single-CTA kernel only): the work of all jobs is laid out as one sequence of
// units - a 2x2 tile block of one segment over its whole K - measured in tile products, and cut into ONE
// CONTIGUOUS SHARE PER CTA: a CTA's piece of a unit is a single item, however long, so its accumulators
// stay in TMEM for the whole piece and are drained once (about 5 drains per CTA instead of 20 on
// ResNet-50).  Parity-tested, and measured SLOWER: 3.81 ms against 2.78 ms for the whole ResNet-50 factor
// stage - the units differ too much in speed per tile product (TMA-fed, gathered, fused-finish) for equal
// shares of tile products to take equal time, and contiguous shares give one CTA a run of like units;
// the short chunks of the default schedule even that out.  Kept behind the switch for that record.
void plan_syrk_tc(FactorPlan* plan, int num_sms) {
  static const int64_t kChunk = getenv("KFB_SYRK_CHUNK") ? atoll(getenv("KFB_SYRK_CHUNK")) : 8192;     // rows of X per item
  const bool stream_k = getenv("KFB_SYRK_STREAMK") && getenv("KFB_SYRK_STREAMK")[0] == '1';    // opt-in, see above
  const int pair_min_tiles = syrk_pair_mode();
  std::vector<TcItem> pair_items, single_items;
  for (size_t ji = 0; ji < plan->jobs.size(); ++ji) {
    const DevJob& j = plan->jobs[ji];
    if (!j.use_tc) continue;
    const int T = ceil_div(j.d, TM);          // 128-wide tiles per side
    const int TB = ceil_div(T, 2);            // 256-wide blocks per side
    const bool pair = pair_min_tiles > 0 && T >= pair_min_tiles;
    std::vector<TcItem>& out = pair ? pair_items : single_items;
    for (int si = 0; si < j.n_segments; ++si) {
      const DevSegment& s = plan->segs[j.seg_start + si];
      // TMA segments iterate the per-image padded matrix: nb blocks of 32 rows per image
      const int64_t K = s.tma >= 0 ? (s.K / ((int64_t)s.oh * s.ow)) * s.nb * BK : s.K;
      const bool whole = j.fused || (stream_k && !pair);      // one unit per block; cut later
      const int64_t nch = whole ? 1 : std::max<int64_t>(1, ceil_div64(K, kChunk));
      const int64_t len = ceil_div64(ceil_div64(K, nch), BK) * BK;
      if (j.la) plan->jobs[ji].la_items += (int)ceil_div64(K, len);     // arrivals every block waits for
      for (int bi = 0; bi < TB; ++bi)
        for (int bj = bi; bj < TB; ++bj)
          for (int64_t k = 0; k < K; k += len) {
            TcItem it;
            it.job = (int)ji; it.seg = j.seg_start + si;
            it.a0 = 2 * bi; it.b0 = 2 * bj;
            it.mi = std::min(2, T - it.a0); it.nj = std::min(2, T - it.b0);
            it.k_begin = (uint32_t)k; it.k_end = (uint32_t)std::min<int64_t>(K, k + len);
            out.push_back(it);
          }
    }
  }
  auto units_of = [](const TcItem& a) { return (int64_t)(a.a0 == a.b0 ? (a.mi * (a.mi + 1)) / 2 : a.mi * a.nj); };
  auto cost = [&](const TcItem& a) { return (int64_t)(a.k_end - a.k_begin) * units_of(a); };
  auto by_cost = [&](const TcItem& a, const TcItem& b) { return cost(a) > cost(b); };
  auto by_len = [](const TcItem& a, const TcItem& b) { return a.k_end - a.k_begin > b.k_end - b.k_begin; };
  std::stable_sort(pair_items.begin(), pair_items.end(), by_len);      // a pair always does the full block
  plan->n_pair_items = (int)pair_items.size();
  plan->tc_items = pair_items;
  plan->cta_first.clear();
  if (!stream_k) {
    std::stable_sort(single_items.begin(), single_items.end(), by_cost);
    plan->tc_items.insert(plan->tc_items.end(), single_items.begin(), single_items.end());
    return;
  }
  // ---- cut the unit sequence into num_sms shares of equal cost
  const int G = std::max(1, num_sms);
  int64_t total = 0;
  for (const TcItem& u : single_items) total += cost(u);
  // every item also pays a fixed price (pipeline refill + drain), charged as rows of a full block
  const int64_t kItemOverhead = 4 * 256;
  total += (int64_t)single_items.size() * kItemOverhead;
  const int64_t share = std::max<int64_t>(1, ceil_div64(total, G));
  const int64_t kMinPiece = 8 * BK;                     // never leave or take a sliver of a unit
  std::vector<TcItem> cut;
  plan->cta_first.assign(1, 0);
  int64_t used = 0;
  auto close_cta = [&]() {
    if ((int)plan->cta_first.size() < G) { plan->cta_first.push_back((int)cut.size()); used = 0; }   // the last CTA takes what is left
  };
  for (const TcItem& u : single_items) {
    const int64_t per_row = units_of(u);
    const bool fused = plan->jobs[u.job].fused != 0;
    int64_t k = u.k_begin;
    while (k < (int64_t)u.k_end) {
      const int64_t left = (int64_t)u.k_end - k;
      int64_t take = left;
      if (!fused) {
        int64_t room = (share - used - kItemOverhead) / per_row;
        room = room / BK * BK;
        const bool last_cta = (int)plan->cta_first.size() >= G;
        if (!last_cta && room < std::min(left, kMinPiece) && used > 0) { close_cta(); continue; }
        if (!last_cta && room < left) {
          take = std::max(room, std::min(left, kMinPiece));
          if (left - take < kMinPiece) take = left;      // do not leave a sliver behind
        }
      }
      TcItem it = u;
      it.k_begin = (uint32_t)k; it.k_end = (uint32_t)(k + take);
      cut.push_back(it);
      used += take * per_row + kItemOverhead;
      k += take;
      if (used >= share) close_cta();
    }
  }
  while ((int)plan->cta_first.size() < G + 1) plan->cta_first.push_back((int)cut.size());
  plan->cta_first.back() = (int)cut.size();
  // positions in tc_items are offset by the pair items, which come first
  for (int& f : plan->cta_first) f += plan->n_pair_items;
  plan->tc_items.insert(plan->tc_items.end(), cut.begin(), cut.end());
}

int launch_syrk_tc(const FactorPlan& plan, void* workspace, Context* ctx, float alpha, cudaStream_t stream) {
  if (plan.tc_items.empty()) return KFB_OK;
  static bool attr_set[64] = {};
  if (ctx->device >= 0 && ctx->device < 64 && !attr_set[ctx->device]) {
    KFB_CUDA(cudaFuncSetAttribute(syrk_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<false>::kSmemBytes));
    KFB_CUDA(cudaFuncSetAttribute(syrk_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<true>::kSmemBytes));
    attr_set[ctx->device] = true;
  }
  char* ws = (char*)workspace;
  const DevJob* jobs = (const DevJob*)(ws + plan.off_jobs);
  const DevSegment* segs = (const DevSegment*)(ws + plan.off_segs);
  const TcItem* items = (const TcItem*)(ws + plan.off_tc_items);
  const CUtensorMap* tmaps = (const CUtensorMap*)(ws + plan.off_tmaps);
  const int* cta_first = plan.cta_first.empty() ? nullptr : (const int*)(ws + plan.off_cta_first);
  const int n_pair = plan.n_pair_items, n_single = (int)plan.tc_items.size() - n_pair;
  if (n_pair > 0) {
    KFB_REQUIRE(ctx->num_sms >= 2, "the CTA-pair SYRK needs at least two SMs");
    const int clusters = std::min(n_pair, ctx->num_sms / 2);     // one item per cluster of 2 CTAs
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters);
    cfg.blockDim = dim3(Cfg<true>::kThreads);
    cfg.dynamicSmemBytes = Cfg<true>::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    KFB_CUDA(cudaLaunchKernelEx(&cfg, syrk_tc_kernel<true>, jobs, segs, items, n_pair, alpha, tmaps, (const int*)nullptr, (int*)nullptr, 0));
    count_launch();
  }
  if (n_single > 0) {
    if (cta_first) {
      // stream-K shares: one CTA per share (cta_first indexes the whole item array)
      const int grid = (int)plan.cta_first.size() - 1;
      syrk_tc_kernel<false><<<grid, Cfg<false>::kThreads, Cfg<false>::kSmemBytes, stream>>>(jobs, segs, items, n_pair + n_single, alpha,
                                                                                           tmaps, cta_first, (int*)nullptr, 0);
    } else {
      const int grid = std::min(n_single, ctx->num_sms);
      // KFB_SYRK_DYNAMIC=1: CTAs claim items from a shared counter instead of the static deal (opt-in: parity-
      // tested, measured equal on ResNet-50 - 2.68-2.74 ms either way - because the early CTAs' idle time is
      // already spent finishing completed blocks)
      const bool dynamic_items = getenv("KFB_SYRK_DYNAMIC") && atoi(getenv("KFB_SYRK_DYNAMIC")) != 0;
      int* la_ctl = plan.la_ctl ? (int*)((float*)(ws + plan.off_scratch) + plan.la_ctl_off) : nullptr;
      syrk_tc_kernel<false><<<grid, Cfg<false>::kThreads, Cfg<false>::kSmemBytes, stream>>>(jobs, segs, items + n_pair, n_single, alpha,
                                                                                           tmaps, (const int*)nullptr, la_ctl, dynamic_items ? 1 : 0);
    }
    KFB_CHECK_LAUNCH();
  }
  return KFB_OK;
}

// (Re-)encode the TMA descriptors of the segments whose capture moved since the last call:
// 3-D map (pixel, channel, image) over the NCHW tensor, box 32 x 128 x 1, 128-byte swizzle - the
// same shared-memory image the gathering producers write.
int refresh_tmaps(FactorPlan* plan, Context* ctx) {
  typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  for (const DevSegment& s : plan->segs) {
    if (s.tma < 0 || plan->tmap_ptr[s.tma] == s.data) continue;
    KFB_REQUIRE(ctx->encode_tiled, "TMA descriptor encoder is not available");
    const int64_t S = (int64_t)s.oh * s.ow, N = s.K / S;
    cuuint64_t dims[3] = {(cuuint64_t)S, (cuuint64_t)s.c, (cuuint64_t)N};
    cuuint64_t strides[2] = {(cuuint64_t)S * sizeof(float), (cuuint64_t)S * s.c * sizeof(float)};
    cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)TM, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = ((EncodeTiledFn)ctx->encode_tiled)(&plan->tmaps[s.tma], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)s.data,
                                                    dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
      KFB_FAIL(KFB_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for an NCHW capture %p [%lld, %d, %lld]", (int)r, s.data,
               (long long)N, s.c, (long long)S);
    plan->tmap_ptr[s.tma] = s.data;
  }
  return KFB_OK;
}

int launch_syrk(const FactorPlan& plan, void* workspace, Context* ctx, float alpha, cudaStream_t stream) {
  int rc = launch_syrk_tc(plan, workspace, ctx, alpha, stream);
  if (rc) return rc;
  return launch_syrk_simt(plan, workspace, stream);
}

}  // namespace kfb
