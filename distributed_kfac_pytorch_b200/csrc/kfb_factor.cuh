// Device-side descriptors and the host plan of one kfb_factor_update call.
#pragma once
#include <vector>

#include <cuda.h>        // CUtensorMap (type only)
#include <cuda_fp16.h>

#include "kfb_common.cuh"

namespace kfb {

struct DevSegment {
  const void* data;
  int dtype;             // KFB_F32 / KFB_F16 / KFB_BF16
  int kind, has_bias;
  int c;                 // features without the bias column
  int h, w;              // NCHW: spatial extent; IM2COL: input extent
  int kh, kw, sh, sw, ph, pw;
  int oh, ow;            // rows per image = oh*ow
  int64_t ld;            // ROWS pitch
  int64_t K;             // rows of X
  float weight;
  const float* dev_scale;   // NULL or device multiplier of weight; 0 drops the segment (kfb_capture_filter)
  int n_chunks;          // split-K chunks of this segment
  int chunk_start;       // first chunk index inside the job
  int64_t chunk_len;
  uint32_t mul_s, sh_s;  // x / (oh*ow) = umulhi(x, mul_s) >> sh_s for x < 2^31 (mul_s = 0: divisor 1)
  uint32_t mul_w, sh_w;  // x / ow likewise
  int nomask;            // IM2COL read from a staging copy (PadJob): h x w is the staged plane, ow the padded
                         // row length, K counts padded pixels; every tap of every X row is in range
  int ow_valid;          // nomask: output pixels per row that exist (bias feature = 1 only there)
  int tma;               // index of this segment's TMA descriptor, -1: gathered by the producer warps
  int nb;                // TMA segments: 32-row blocks per image (the last one zero-filled past oh*ow)
  int fmt;               // 0: fp32 operands, kind::tf32.  KFB_F16 / KFB_BF16: the operand is 16-bit and TMA-fed; it is
                         // DESCRIBED as an fp32 matrix of half as many columns (two X rows per 4-byte word: same bytes,
                         // same 128-byte tile rows, K counted in words) and only the MMA differs: kind::f16 reads every
                         // 128-byte row as 64 K-elements - exact products, half the bytes, twice the MMA rate
};

struct DevJob {
  int d;
  int seg_start, n_segments;
  float* state;
  int64_t ld_state;
  int seed_identity, write_state;
  float* tri_out;
  const float* dev_rescale;  // NULL or device float: >= 0 multiplies F, < 0 leaves the running average untouched
  float* scratch;        // d x d accumulator (upper triangle used), pitch ld_scratch
  int64_t ld_scratch;
  size_t scratch_off;    // host only: float offset into the scratch region
  int T, n_pairs;        // 64-wide tiles per side, upper-triangle tile pairs
  int n_chunks;          // split-K chunks over all segments
  int T32;               // 32-wide blocks per side for finalize
  int use_tc;            // 1: handled by the tcgen05 SYRK
  int native;            // 1: the scratch accumulator is laid out the way the tensor-core epilogue holds it (a thread
                         // owns an accumulator row): [128x128 tile (row-major grid of T128^2)][32-row quadrant q]
                         // [32-col chunk c][4-col group g][row lane][4], so one red.v4 of a warp is 512 contiguous
                         // bytes instead of 32 lines; finalize reads whole 32x32 blocks = one 4 KB chunk
  int T128;              // 128-wide tiles per side
  int fused;             // 1: one K chunk, one segment: the SYRK epilogue applies the running average
                         //    itself and writes the state - no scratch, no reds, no finalize pass
  int la;                // 1: split-K job finished by the LAST item of every 256 x 256 block to arrive (per-block
                         //    counter): that item reads the block's sum back from the scratch, re-zeroes scratch and
                         //    counter, and applies the running average as a fused job does - no finalize pass
  int la_items;          // items (K chunks over all segments) per block
  int* arrive;           // [ceil(T128/2)^2] arrival counters, zero between updates
  size_t arrive_off;     // host only: float offset of the counters inside the scratch region
};

// One captured element, widened to fp32 (16-bit captures come from AMP autocast).
__device__ __forceinline__ float widen16(unsigned short u, int dtype) {
  return dtype == KFB_F16 ? __half2float(__ushort_as_half(u)) : __uint_as_float((uint32_t)u << 16);
}
__device__ __forceinline__ float load_elem(const void* base, int dtype, int64_t idx) {
  if (dtype == KFB_F32) return __ldg(reinterpret_cast<const float*>(base) + idx);
  return widen16(__ldg(reinterpret_cast<const unsigned short*>(base) + idx), dtype);
}

// Staging copy of one im2col capture for the tensor-core SYRK (fp32, whatever the capture's type).
// For every (image, channel) plane of the capture it holds kw planes of hp x wp floats, one per
// kernel column kx, with the conv's x-stride already applied:
//   dst[(plane*kw + kx)][y][x] = src[plane][y*sy - pad_h][x*sx + kx - pad_w]   for x < ow_valid,
//   0 where that is outside the source image, and 0 for the pitch padding x >= ow_valid.
// kxk convs: sy = 1, sx = stride, hp = (oh-1)*stride_h + kh, wp = ow rounded up to 4.  Four
// consecutive output pixels of one (ky,kx) tap are then 16 contiguous, 16-byte aligned bytes, so the
// gather moves 16 bytes per lane and needs no border tests.
// 1x1 strided convs: kw = 1, sy/sx = stride, hp x wp = oh x ow: the dense sub-sampled image, which
// then is an ordinary NCHW operand.
// Variant 4 - NCHW operands TMA cannot take as they are (16-bit captures, image rows that are not
// 16-byte multiples such as 7x7 maps, misaligned views): fp32 copy with images and channels swapped,
//   dst[c][n*wp + s] = src[n][c][s]  (s < ow_valid = h*w of the capture; wp = that rounded up to 4),
// which is a plain 2-D [channels][rows of X] matrix for TMA, with no per-image block padding.
struct PadJob {
  const void* src;
  float* dst;
  int dtype;
  int h, w;              // source image
  int hp, wp;            // staged plane (wp = row pitch)
  int kw, ow_valid;
  int sy, sx, pad_h, pad_w;
  uint32_t mul_wp, sh_wp, mul_hp, sh_hp, mul_kw, sh_kw;   // x / wp, x / hp, x / kw by multiply-high (magic_div)
  uint32_t mul_wq, sh_wq;                                  // x / (wp/4)
  int variant;           // 0: one thread per 4 output floats; 1: (sx,kw) = (1,3), 2: (2,3), 3: (2,7): one thread
                         // reads a source row segment once and writes the 4-pixel group of all kw planes;
                         // 4: NCHW capture -> [channels][images * wp] (see below)
                         // 5: row-major capture -> [features (+ ones)][rows rounded up to 32] (pad_rows)
  int channels;          // variant 4
  int out16;             // variants 4, 5: the copy keeps the capture's 16-bit type (DevSegment::fmt) instead of widening
  int64_t quads;         // variants 1-3: planes * hp * wp/4 threads; variant 5: 32-row tiles along K
  int64_t total;         // planes * kw * hp * wp (< 2^31)
};

// float index of element (i, j) in a DevJob::native scratch
__host__ __device__ __forceinline__ int64_t native_index(int T128, int i, int j) {
  const int64_t chunk = ((int64_t)((i >> 7) * T128 + (j >> 7)) * 16 + ((i >> 5) & 3) * 4 + ((j >> 5) & 3));
  return chunk * 1024 + ((((j >> 2) & 7) * 32 + (i & 31)) << 2) + (j & 3);
}

struct TcItem {            // one unit of tensor-core SYRK work: a 2x2 block of 128-wide tiles
  int job, seg;
  int a0, b0;              // first row tile / first column tile (a0 <= b0; equal on the diagonal)
  int mi, nj;              // tiles in the block's rows / columns (1 or 2)
  uint32_t k_begin, k_end; // rows of X (TMA segments: rows of the per-image padded matrix, nb*32 per image)
};

struct FactorPlan {
  std::vector<TcItem> tc_items;          // [0, n_pair_items): CTA-pair kernel; the rest: single-CTA kernel
  int n_pair_items = 0;
  size_t off_tc_items = 0;
  std::vector<int> cta_first;            // single-CTA kernel, stream-K shares: items [cta_first[c], cta_first[c+1]) belong to CTA c
  size_t off_cta_first = 0;
  std::vector<DevJob> jobs;
  std::vector<DevSegment> segs;
  std::vector<int> item_start, fin_start;
  std::vector<PadJob> pads;              // staging copies, one per staged segment
  std::vector<int> seg_pad;              // per segment: index into pads or -1
  std::vector<size_t> pad_off;           // float offset of each copy inside the staging region
  std::vector<int> pad_start;            // first block of each PadJob (+ total)
  size_t stage_floats = 0, off_stage = 0, off_pads = 0, off_pad_start = 0;
  std::vector<CUtensorMap> tmaps;        // one per TMA-fed segment
  std::vector<const void*> tmap_ptr;     // the address each descriptor was encoded for
  size_t off_tmaps = 0;
  int n_items = 0, n_fin = 0;
  size_t scratch_floats = 0;             // [accumulators cleared by every call][last-arriver accumulators + counters]
  size_t la_begin = 0;                   // float offset of the second part: left zero by the kernel itself
  size_t la_ctl_off = 0;                 // float offset of the completed-block queue inside the scratch region
  int la_blocks = 0;                     // 256 x 256 blocks of all last-arriver jobs
  bool la_ctl = false;                   // the control block exists: dynamic item hand-out + completed-block queue
  uint64_t generation = 0;               // bumped whenever the plan is rebuilt (the layout may have moved)
  size_t off_jobs = 0, off_segs = 0, off_items = 0, off_fin = 0, off_scratch = 0, total_bytes = 0;
};

KFB_HIDDEN int plan_factor(const kfb_factor_job* jobs, int n_jobs, int num_sms, bool allow_tc, FactorPlan* plan);
KFB_HIDDEN void plan_syrk_tc(FactorPlan* plan, int num_sms);
KFB_HIDDEN int upload_plan(FactorPlan* plan, void* workspace, cudaStream_t stream, bool la_clean = false);
KFB_HIDDEN int refresh_tmaps(FactorPlan* plan, Context* ctx);
KFB_HIDDEN void bind_workspace(FactorPlan* plan, void* workspace);
KFB_HIDDEN int launch_pad(const FactorPlan& plan, void* workspace, cudaStream_t stream);
KFB_HIDDEN int launch_syrk_simt(const FactorPlan& plan, void* workspace, cudaStream_t stream);
KFB_HIDDEN int launch_syrk(const FactorPlan& plan, void* workspace, Context* ctx, float alpha, cudaStream_t stream);
constexpr int kLaCtlInts = 64;            // control words [pushed, popped, CTAs gone, items handed out] in front of the completed-block queue
constexpr int64_t kFusedMaxRows = 4096;   // jobs with at most this many rows of X are never split along K
KFB_HIDDEN int factor_update_with(const kfb_factor_job* jobs, int n_jobs, float alpha, Context* ctx, void* workspace,
                                  size_t workspace_bytes, const char* what, cudaStream_t stream);
KFB_HIDDEN int dedicate_workspace(void* workspace, size_t bytes, bool on);
KFB_HIDDEN size_t factor_workspace_bytes(const kfb_factor_job* jobs, int n_jobs, Context* ctx);
KFB_HIDDEN int launch_finalize(const FactorPlan& plan, void* workspace, float alpha, cudaStream_t stream);

}  // namespace kfb
