// Factor computation: symmetric rank-k update over implicitly-unfolded activations, split-K
// accumulation into caller scratch, and a finalize pass that fuses the scale, the running
// average, the mirror and the packed-triangle export.
//
// This file holds the host planning, the SIMT fp32 SYRK (exact-fp32 path, also the path for
// shapes the tensor-core kernel does not take) and the finalize / pack / unpack kernels.
// The tcgen05 SYRK lives in kfb_syrk_tc.cu and writes the same scratch convention.
#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <vector>

#include "kfb_factor.cuh"

namespace kfb {

// ============================================================================================
// SIMT SYRK: one CTA = one 64x64 upper-triangle tile pair x one K chunk of one segment
// ============================================================================================
namespace {

constexpr int TS = 64;     // tile side
constexpr int BK = 32;     // K step
constexpr int NT = 256;    // threads

struct IDecode {           // decode of a feature index for IM2COL
  int off;                 // ci*h*w + ky*w + kx
  int ky, kx;
  int valid;               // 0: index >= d (padding), 1: data, 2: bias column
};

__device__ __forceinline__ IDecode decode_feature(const DevSegment& s, int d, int i) {
  IDecode r;
  r.off = 0; r.ky = 0; r.kx = 0;
  if (i >= d) { r.valid = 0; return r; }
  if (s.has_bias && i == d - 1) { r.valid = 2; return r; }
  r.valid = 1;
  if (s.kind == KFB_SEG_IM2COL) {
    int khw = s.kh * s.kw;
    int ci = i / khw;
    int q = i - ci * khw;
    r.ky = q / s.kw;
    r.kx = q - r.ky * s.kw;
    r.off = (ci * s.h + r.ky) * s.w + r.kx;
  } else if (s.kind == KFB_SEG_NCHW) {
    r.off = i * s.h * s.w;
  } else {
    r.off = i;
  }
  return r;
}

// Loads a BK x TS operand tile: dst[kk][ii] = X[k0+kk][i0+ii] (0 outside).
__device__ __forceinline__ void load_tile(const DevSegment& s, int d, int64_t k0, int64_t k_end,
                                          const IDecode* dec /*8 per thread, K-major map*/,
                                          int i0, float (*dst)[TS + 1]) {
  const int t = threadIdx.x;
  if (s.kind == KFB_SEG_ROWS) {
    // features are contiguous: lanes run along ii
    const int ii = t & (TS - 1);
    const int i = i0 + ii;
    const bool bias = s.has_bias && i == d - 1;
    const bool in = i < d;
#pragma unroll
    for (int j = 0; j < BK / (NT / TS); ++j) {
      int kk = (t >> 6) + j * (NT / TS);
      int64_t k = k0 + kk;
      float v = 0.f;
      if (in && k < k_end) v = bias ? 1.f : load_elem(s.data, s.dtype, k * s.ld + i);
      dst[kk][ii] = v;
    }
  } else {
    // rows (k) are contiguous in memory: lanes run along kk
    const int kk = t & (BK - 1);
    const int64_t k = k0 + kk;
    const bool kin = k < k_end;
    const int S = s.oh * s.ow;
    int64_t base = 0;
    int y0 = 0, x0 = 0;
    if (kin) {
      int64_t n = k / S;
      int r = (int)(k - n * S);
      if (s.kind == KFB_SEG_NCHW) {
        base = n * (int64_t)s.c * S + r;
      } else {
        int oy = r / s.ow;
        int ox = r - oy * s.ow;
        y0 = oy * s.sh - s.ph;
        x0 = ox * s.sw - s.pw;
        base = n * (int64_t)s.c * s.h * s.w + (int64_t)y0 * s.w + x0;
      }
    }
#pragma unroll
    for (int j = 0; j < TS / (NT / BK); ++j) {
      int ii = (t >> 5) + j * (NT / BK);
      const IDecode& e = dec[j];
      float v = 0.f;
      if (kin && e.valid) {
        if (e.valid == 2) {
          v = 1.f;
        } else if (s.kind == KFB_SEG_NCHW) {
          v = load_elem(s.data, s.dtype, base + e.off);
        } else {
          int y = y0 + e.ky, x = x0 + e.kx;
          if (y >= 0 && y < s.h && x >= 0 && x < s.w) v = load_elem(s.data, s.dtype, base + e.off);
        }
      }
      dst[kk][ii] = v;
    }
  }
}

__global__ void __launch_bounds__(NT) syrk_simt_kernel(const DevJob* __restrict__ jobs,
                                                      const DevSegment* __restrict__ segs,
                                                      const int* __restrict__ item_start,
                                                      int n_jobs) {
  __shared__ float As[BK][TS + 1];
  __shared__ float Bs[BK][TS + 1];

  // ---- which work item am I
  int lo = 0, hi = n_jobs;
  const int b = blockIdx.x;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (item_start[mid] <= b) lo = mid; else hi = mid;
  }
  const DevJob job = jobs[lo];
  int local = b - item_start[lo];
  int pair = local % job.n_pairs;
  int chunk = local / job.n_pairs;
  // pair -> (ti, tj), ti <= tj, row-major over the upper triangle
  int ti = 0, rem = pair, rowlen = job.T;
  while (rem >= rowlen) { rem -= rowlen; --rowlen; ++ti; }
  int tj = ti + rem;
  // chunk -> segment
  int sidx = job.seg_start;
  while (chunk >= segs[sidx].chunk_start + segs[sidx].n_chunks) ++sidx;
  const DevSegment s = segs[sidx];
  chunk -= s.chunk_start;
  const int64_t k_begin = (int64_t)chunk * s.chunk_len;
  const int64_t k_end = min(s.K, k_begin + s.chunk_len);
  const int d = job.d;
  const int i0 = ti * TS, j0 = tj * TS;
  const bool diag = ti == tj;

  IDecode decA[8], decB[8];
  if (s.kind != KFB_SEG_ROWS) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int ii = (threadIdx.x >> 5) + j * 8;
      decA[j] = decode_feature(s, d, i0 + ii);
      decB[j] = decode_feature(s, d, j0 + ii);
    }
  }

  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4];
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[p][q] = 0.f;

  float (*Bt)[TS + 1] = diag ? As : Bs;
  for (int64_t k0 = k_begin; k0 < k_end; k0 += BK) {
    load_tile(s, d, k0, k_end, decA, i0, As);
    if (!diag) load_tile(s, d, k0, k_end, decB, j0, Bs);
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], bb[4];
#pragma unroll
      for (int p = 0; p < 4; ++p) a[p] = As[kk][ty * 4 + p];
#pragma unroll
      for (int q = 0; q < 4; ++q) bb[q] = Bt[kk][tx * 4 + q];
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[p][q] = fmaf(a[p], bb[q], acc[p][q]);
    }
    __syncthreads();
  }

  float wgt = s.weight;
  if (s.dev_scale) {
    const float m = __ldg(s.dev_scale);
    if (m == 0.f) return;              // dropped capture: its sums may be inf / nan
    wgt *= m;
  }
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    int i = i0 + ty * 4 + p;
    if (i >= d) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      int j = j0 + tx * 4 + q;
      if (j >= d || j < i) continue;
      atomicAdd(job.scratch + (int64_t)i * job.ld_scratch + j, acc[p][q] * wgt);
    }
  }
}

// ============================================================================================
// finalize: new = (old * alpha/(1-alpha) + F) * (1-alpha); mirror; optional packed export
// ============================================================================================
__global__ void __launch_bounds__(256) finalize_kernel(const DevJob* __restrict__ jobs,
                                                      const int* __restrict__ fin_start,
                                                      int n_jobs, float alpha) {
  __shared__ float tile[32][33];
  int lo = 0, hi = n_jobs;
  const int b = blockIdx.x;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (fin_start[mid] <= b) lo = mid; else hi = mid;
  }
  const DevJob job = jobs[lo];
  int rem = b - fin_start[lo], rowlen = job.T32, bi = 0;
  while (rem >= rowlen) { rem -= rowlen; --rowlen; ++bi; }
  const int bj = bi + rem;
  const int d = job.d;
  const float ratio = alpha / (1.f - alpha);
  const float one_m = 1.f - alpha;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  const float resc = job.dev_rescale ? __ldg(job.dev_rescale) : 1.f;   // < 0: no capture survived, keep the old average

  // pass 1: rows of the (bi, bj) block; value valid where j >= i
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    int li = ty + r * 8;
    int i = bi * 32 + li, j = bj * 32 + tx;
    float v = 0.f;
    if (i < d && j < d && j >= i) {
      float f = job.scratch[job.native ? native_index(job.T128, i, j) : (int64_t)i * job.ld_scratch + j];
      float old = job.seed_identity ? (i == j ? 1.f : 0.f) : job.state[(int64_t)i * job.ld_state + j];
      v = resc < 0.f ? old : (old * ratio + f * resc) * one_m;
      if (job.tri_out) job.tri_out[(int64_t)i * d - ((int64_t)i * (i - 1)) / 2 + (j - i)] = v;
    }
    tile[li][tx] = v;
  }
  __syncthreads();
  if (!job.write_state) return;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    int li = ty + r * 8;
    int i = bi * 32 + li, j = bj * 32 + tx;
    if (i < d && j < d) {
      // diagonal block: the strictly-lower part comes from the transposed element
      float v = (j >= i) ? tile[li][tx] : tile[tx][li];
      job.state[(int64_t)i * job.ld_state + j] = v;
    }
    if (bi != bj) {   // mirrored block (bj, bi): element (bj*32+li, bi*32+tx) = tile[tx][li]
      int mi = bj * 32 + li, mj = bi * 32 + tx;
      if (mi < d && mj < d) job.state[(int64_t)mi * job.ld_state + mj] = tile[tx][li];
    }
  }
}

// ============================================================================================
// packed upper triangle <-> full symmetric square
// ============================================================================================
struct TriJob {
  float* tri;
  float* state;
  int64_t ld;
  int d, T32;
};
constexpr int kTriBatch = 512;
typedef Batch<TriJob, kTriBatch> TriBatch;

template <bool kUnpack>
__global__ void __launch_bounds__(256) tri_kernel(const __grid_constant__ TriBatch batch, float scale) {
  __shared__ float tile[32][33];
  const int jidx = find_job(batch, blockIdx.x);
  const TriJob job = batch.jobs[jidx];
  int rem = blockIdx.x - batch.start[jidx], rowlen = job.T32, bi = 0;
  while (rem >= rowlen) { rem -= rowlen; --rowlen; ++bi; }
  const int bj = bi + rem;
  const int d = job.d;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    int li = ty + r * 8;
    int i = bi * 32 + li, j = bj * 32 + tx;
    float v = 0.f;
    if (i < d && j < d && j >= i) {
      int64_t o = (int64_t)i * d - ((int64_t)i * (i - 1)) / 2 + (j - i);
      if (kUnpack) {
        v = job.tri[o] * scale;
      } else {
        job.tri[o] = job.state[(int64_t)i * job.ld + j];
      }
    }
    tile[li][tx] = v;
  }
  if (!kUnpack) return;
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    int li = ty + r * 8;
    int i = bi * 32 + li, j = bj * 32 + tx;
    if (i < d && j < d) job.state[(int64_t)i * job.ld + j] = (j >= i) ? tile[li][tx] : tile[tx][li];
    if (bi != bj) {
      int mi = bj * 32 + li, mj = bi * 32 + tx;
      if (mi < d && mj < d) job.state[(int64_t)mi * job.ld + mj] = tile[tx][li];
    }
  }
}

// Staging copies for the tensor-core SYRK (PadJob): every thread writes four consecutive floats of
// one copy.  HBM-bound: reads the capture once, writes 1.0-1.2x its size.
// The staging copies exist only for the tensor-core SYRK, whose kind::tf32 MMAs read the top 19 bits of
// an fp32 operand, i.e. TRUNCATE it.  Truncation is biased: every x*x term of a diagonal entry comes
// out low by about 2^-10 of itself, which is most of the factor error budget (3.7e-4 of the 1e-3
// tolerance, normalised, on the 4608-dim ResNet-50 factor).  Rounding to nearest here costs nothing (the
// value passes through a register anyway) and makes the operand error unbiased, so it averages out
// over K.  16-bit captures are exact in TF32.
template <typename T>
__device__ __forceinline__ float pad_load(const T* p, int dtype) {
  if (sizeof(T) == 4) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(__ldg(reinterpret_cast<const float*>(p))));
    return __uint_as_float(r);
  }
  return widen16(__ldg(reinterpret_cast<const unsigned short*>(p)), dtype);
}

// fp32 value that came from a 16-bit capture -> its 16 bits again (exact)
__device__ __forceinline__ uint32_t narrow16(float v, int dtype) {
  return dtype == KFB_F16 ? (uint32_t)__half_as_ushort(__float2half_rn(v)) : (__float_as_uint(v) >> 16);
}
// four consecutive elements of a 16-bit staging copy (PadJob::out16)
__device__ __forceinline__ void store4_16(float* dst, size_t e, const float* v, int dtype) {
  *reinterpret_cast<uint2*>(reinterpret_cast<unsigned short*>(dst) + e) =
      make_uint2(narrow16(v[0], dtype) | (narrow16(v[1], dtype) << 16), narrow16(v[2], dtype) | (narrow16(v[3], dtype) << 16));
}

template <typename T>
__device__ __forceinline__ void pad_quad(const PadJob& p, uint32_t e0) {
  // (plane', y, x) of the first of this thread's four consecutive floats; plane' = plane*kw + kx
  const uint32_t row = p.mul_wp ? (__umulhi(e0, p.mul_wp) >> p.sh_wp) : e0;
  int x = (int)(e0 - row * (uint32_t)p.wp);
  uint32_t pk = p.mul_hp ? (__umulhi(row, p.mul_hp) >> p.sh_hp) : row;
  int y = (int)(row - pk * (uint32_t)p.hp);
  uint32_t plane = p.mul_kw ? (__umulhi(pk, p.mul_kw) >> p.sh_kw) : pk;
  int kx = (int)(pk - plane * (uint32_t)p.kw);
  const size_t plane_elems = (size_t)p.h * p.w;
  const T* img = reinterpret_cast<const T*>(p.src) + (size_t)plane * plane_elems;
  const uint32_t total = (uint32_t)p.total;
  float v[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int sy = y * p.sy - p.pad_h, sx = x * p.sx + kx - p.pad_w;
    v[i] = 0.f;
    if (e0 + i < total && x < p.ow_valid && (unsigned)sy < (unsigned)p.h && (unsigned)sx < (unsigned)p.w)
      v[i] = pad_load(img + sy * p.w + sx, p.dtype);
    if (++x == p.wp) {
      x = 0;
      if (++y == p.hp) { y = 0; if (++kx == p.kw) { kx = 0; img += plane_elems; } }
    }
  }
  if (sizeof(T) == 2 && p.out16) {        // total is a multiple of 8 there
    store4_16(p.dst, e0, v, p.dtype);
  } else if (e0 + 3 < total) {
    *reinterpret_cast<float4*>(p.dst + e0) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
    for (int i = 0; i < 4 && e0 + i < total; ++i) p.dst[e0 + i] = v[i];
  }
}

// The common conv shapes: one thread per (plane, row, 4-pixel group) loads the source row segment
// that the group's kw taps touch ONCE and writes the group of all kw planes (3-7x fewer loads and
// index decodes per output float than pad_quad).
template <typename T, int SX, int KW>
__device__ __forceinline__ void pad_group(const PadJob& p, uint32_t q) {
  const uint32_t wq = (uint32_t)p.wp >> 2;
  const uint32_t row = p.mul_wq ? (__umulhi(q, p.mul_wq) >> p.sh_wq) : q;
  const int x0 = (int)(q - row * wq) * 4;
  const uint32_t plane = p.mul_hp ? (__umulhi(row, p.mul_hp) >> p.sh_hp) : row;
  const int y = (int)(row - plane * (uint32_t)p.hp);
  const int sy = y - p.pad_h;                       // these variants have sy = 1
  constexpr int kSpan = 3 * SX + KW;
  float v[kSpan];
#pragma unroll
  for (int j = 0; j < kSpan; ++j) v[j] = 0.f;
  if ((unsigned)sy < (unsigned)p.h) {
    const T* srow = reinterpret_cast<const T*>(p.src) + ((size_t)plane * p.h + sy) * p.w;
    const int s0 = x0 * SX - p.pad_w;
#pragma unroll
    for (int j = 0; j < kSpan; ++j)
      if ((unsigned)(s0 + j) < (unsigned)p.w) v[j] = pad_load(srow + s0 + j, p.dtype);
  }
  const size_t out_e = ((size_t)plane * KW * p.hp + y) * p.wp + x0;
  float* out = p.dst + out_e;
  const size_t plane_stride = (size_t)p.hp * p.wp;
  const int nvalid = p.ow_valid - x0;               // pixels of this group that exist (may be <= 0)
#pragma unroll
  for (int kx = 0; kx < KW; ++kx) {
    float4 o;
    o.x = nvalid > 0 ? v[kx] : 0.f;
    o.y = nvalid > 1 ? v[SX + kx] : 0.f;
    o.z = nvalid > 2 ? v[2 * SX + kx] : 0.f;
    o.w = nvalid > 3 ? v[3 * SX + kx] : 0.f;
    if (sizeof(T) == 2 && p.out16) {
      const float ov[4] = {o.x, o.y, o.z, o.w};
      store4_16(p.dst, out_e + kx * plane_stride, ov, p.dtype);
    } else {
      *reinterpret_cast<float4*>(out + kx * plane_stride) = o;
    }
  }
}

// Variant 4: dst[c][n*wp + s] = src[n][c][s]; hp = images, so (row = e0 / wp) = c*images + n.
template <typename T>
__device__ __forceinline__ void pad_transpose(const PadJob& p, uint32_t e0) {
  const uint32_t row = p.mul_wp ? (__umulhi(e0, p.mul_wp) >> p.sh_wp) : e0;
  const int x0 = (int)(e0 - row * (uint32_t)p.wp);
  const uint32_t c = p.mul_hp ? (__umulhi(row, p.mul_hp) >> p.sh_hp) : row;
  const uint32_t n = row - c * (uint32_t)p.hp;
  const T* srow = reinterpret_cast<const T*>(p.src) + ((size_t)n * p.channels + c) * p.ow_valid;
  if (sizeof(T) == 2 && p.out16) {        // the copy keeps the 16-bit type: raw bits, zero padding
    const unsigned short* s16 = reinterpret_cast<const unsigned short*>(srow) + x0;
    if (x0 + 3 < p.ow_valid && (((uintptr_t)s16) & 7) == 0) {         // four whole pixels in one 8-byte load
      *reinterpret_cast<uint2*>(reinterpret_cast<unsigned short*>(p.dst) + e0) = __ldg(reinterpret_cast<const uint2*>(s16));
      return;
    }
    unsigned short u[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) u[i] = x0 + i < p.ow_valid ? __ldg(reinterpret_cast<const unsigned short*>(srow) + x0 + i) : (unsigned short)0;
    *reinterpret_cast<uint2*>(reinterpret_cast<unsigned short*>(p.dst) + e0) =
        make_uint2((uint32_t)u[0] | ((uint32_t)u[1] << 16), (uint32_t)u[2] | ((uint32_t)u[3] << 16));
    return;
  }
  float v[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = x0 + i < p.ow_valid ? pad_load(srow + x0 + i, p.dtype) : 0.f;
  *reinterpret_cast<float4*>(p.dst + e0) = make_float4(v[0], v[1], v[2], v[3]);     // wp % 4 == 0
}

// Variant 5: dst[c][k] = src[k][c] for a row-major capture (Linear layers: h = features, w = pitch, hp = rows,
// wp = rows rounded up to 32), plus a row of ones (bias) when channels = h + 1; k >= rows is zero.  One 32 x 32
// tile per block through shared memory: reads run along the features, writes along k.
template <typename T>
__device__ __forceinline__ void pad_rows(const PadJob& p, int b) {
  __shared__ float tile[32][33];
  const int tiles_k = (int)p.quads;
  const int c0 = (b / tiles_k) * 32, k0 = (b % tiles_k) * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const T* src = reinterpret_cast<const T*>(p.src);
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int k = k0 + ty + 8 * r, c = c0 + tx;
    float v = 0.f;
    const bool raw = sizeof(T) == 2 && p.out16;       // 16-bit copy: the tile carries the raw bits
    if (k < p.hp) {
      if (c < p.h) v = raw ? __uint_as_float((uint32_t)__ldg(reinterpret_cast<const unsigned short*>(src) + (size_t)k * p.w + c))
                           : pad_load(src + (size_t)k * p.w + c, p.dtype);
      else if (c < p.channels) v = raw ? __uint_as_float(p.dtype == KFB_F16 ? 0x3C00u : 0x3F80u) : 1.f;    // 1.0
    }
    tile[ty + 8 * r][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int c = c0 + ty + 8 * r;
    if (c >= p.channels) continue;
    if (sizeof(T) == 2 && p.out16)
      reinterpret_cast<unsigned short*>(p.dst)[(size_t)c * p.wp + k0 + tx] = (unsigned short)__float_as_uint(tile[tx][ty + 8 * r]);
    else
      p.dst[(size_t)c * p.wp + k0 + tx] = tile[tx][ty + 8 * r];
  }
}

// Staging copies for the tensor-core SYRK (PadJob).  HBM-bound: reads the capture once, writes kw x
// (about) its size.
__global__ void __launch_bounds__(256) pad_kernel(const PadJob* __restrict__ pads, const int* __restrict__ start, int n_pads) {
  int lo = 0, hi = n_pads;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (start[mid] <= (int)blockIdx.x) lo = mid; else hi = mid;
  }
  const PadJob& p = pads[lo];
  const uint32_t t = (uint32_t)(blockIdx.x - start[lo]) * 256 + threadIdx.x;
  const bool f32 = p.dtype == KFB_F32;
  if (p.variant == 5) {                         // block-uniform
    if (f32) pad_rows<float>(p, blockIdx.x - start[lo]); else pad_rows<unsigned short>(p, blockIdx.x - start[lo]);
    return;
  }
  if (p.variant == 0 || p.variant == 4) {
    const uint32_t e0 = t * 4;
    if (e0 >= (uint32_t)p.total) return;
    if (p.variant == 4) { if (f32) pad_transpose<float>(p, e0); else pad_transpose<unsigned short>(p, e0); }
    else if (f32) pad_quad<float>(p, e0);
    else pad_quad<unsigned short>(p, e0);
    return;
  }
  if (t >= (uint32_t)p.quads) return;
  switch (p.variant) {
    case 1: if (f32) pad_group<float, 1, 3>(p, t); else pad_group<unsigned short, 1, 3>(p, t); break;
    case 2: if (f32) pad_group<float, 2, 3>(p, t); else pad_group<unsigned short, 2, 3>(p, t); break;
    default: if (f32) pad_group<float, 2, 7>(p, t); else pad_group<unsigned short, 2, 7>(p, t); break;
  }
}

int run_tri(const kfb_tri_job* jobs, int n_jobs, float scale, bool unpack, cudaStream_t stream) {
  KFB_REQUIRE(n_jobs >= 0 && (jobs || n_jobs == 0), "bad job array");
  for (int base = 0; base < n_jobs; base += kTriBatch) {
    TriBatch batch;
    batch.n = std::min(kTriBatch, n_jobs - base);
    int blocks = 0;
    for (int i = 0; i < batch.n; ++i) {
      const kfb_tri_job& j = jobs[base + i];
      KFB_REQUIRE(j.tri && j.state && j.d > 0 && j.ld_state >= j.d, "tri job %d is malformed", base + i);
      TriJob& t = batch.jobs[i];
      t.tri = j.tri; t.state = j.state; t.ld = j.ld_state; t.d = j.d; t.T32 = ceil_div(j.d, 32);
      batch.start[i] = blocks;
      blocks += t.T32 * (t.T32 + 1) / 2;
    }
    batch.start[batch.n] = blocks;
    if (unpack) tri_kernel<true><<<blocks, 256, 0, stream>>>(batch, scale);
    else tri_kernel<false><<<blocks, 256, 0, stream>>>(batch, scale);
    KFB_CHECK_LAUNCH();
  }
  return KFB_OK;
}

}  // namespace

// ============================================================================================
// planning (host)
// ============================================================================================
static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
// KFB_SYRK_LAST_ARRIVER=0: split-K jobs go back to scratch + finalize_kernel
static int last_arriver_mode() {
  const char* e = getenv("KFB_SYRK_LAST_ARRIVER");
  return e ? atoi(e) : 1;
}
// KFB_SYRK_ROWS_TC=0: Linear-layer captures go back to the SIMT SYRK
static int rows_tc_mode() {
  const char* e = getenv("KFB_SYRK_ROWS_TC");
  return e ? atoi(e) : 1;
}
// KFB_SYRK_HALF=0: 16-bit captures are widened to fp32 staging copies (kind::tf32) instead of feeding kind::f16
static int half_mode() {
  const char* e = getenv("KFB_SYRK_HALF");
  return e ? atoi(e) : 1;
}
static int stream_k_mode() {
  const char* e = getenv("KFB_SYRK_STREAMK");
  return e ? atoi(e) : 0;
}
static int64_t fused_max_rows() {
  const char* e = getenv("KFB_SYRK_FUSE_MAX_ROWS");
  return e ? atoll(e) : kFusedMaxRows;
}

// Division by an invariant divisor d (1 <= d < 2^31) of any x < 2^31 as a multiply-high and a shift:
// p = 31 + ceil(log2 d), m = floor(2^p / d) + 1 gives floor(x*m / 2^p) = floor(x / d) because
// m*d - 2^p <= d <= 2^(p-31).  d = 1 is flagged by m = 0.
static void magic_div(int64_t d, uint32_t* mul, uint32_t* shift) {
  if (d <= 1) { *mul = 0; *shift = 0; return; }
  int l = 0;
  while (((int64_t)1 << l) < d) ++l;
  const int p = 31 + l;
  *mul = (uint32_t)(((unsigned __int128)1 << p) / (unsigned __int128)d + 1);
  *shift = (uint32_t)(p - 32);
}

int plan_factor(const kfb_factor_job* jobs, int n_jobs, int num_sms, bool allow_tc, FactorPlan* plan) {
  KFB_REQUIRE(n_jobs >= 0 && (jobs || n_jobs == 0), "bad job array");
  plan->jobs.clear(); plan->segs.clear(); plan->item_start.clear(); plan->fin_start.clear();
  plan->tmaps.clear(); plan->tmap_ptr.clear();
  plan->pads.clear(); plan->seg_pad.clear(); plan->pad_off.clear(); plan->pad_start.clear();
  plan->stage_floats = 0;
  plan->scratch_floats = 0;
  plan->la_blocks = 0;
  size_t la_floats = 0;
  int items = 0, fins = 0;
  const int target = std::max(1, num_sms) * 4;
  for (int ji = 0; ji < n_jobs; ++ji) {
    const kfb_factor_job& j = jobs[ji];
    KFB_REQUIRE(j.d > 0 && j.n_segments > 0 && j.segments, "factor job %d: empty", ji);
    KFB_REQUIRE(j.state && j.ld_state >= j.d, "factor job %d: bad state / pitch", ji);
    DevJob dj;
    memset(&dj, 0, sizeof(dj));
    dj.d = j.d;
    dj.seg_start = (int)plan->segs.size();
    dj.n_segments = j.n_segments;
    dj.state = j.state; dj.ld_state = j.ld_state;
    dj.seed_identity = j.seed_identity; dj.write_state = j.write_state; dj.tri_out = j.tri_out;
    dj.dev_rescale = j.dev_rescale;
    dj.ld_scratch = (j.d + 3) / 4 * 4;
    dj.T = ceil_div(j.d, TS);
    dj.n_pairs = dj.T * (dj.T + 1) / 2;
    dj.T32 = ceil_div(j.d, 32);
    int64_t k_total = 0;
    bool tc = allow_tc;
    for (int si = 0; si < j.n_segments; ++si) {
      const kfb_segment& s = j.segments[si];
      KFB_REQUIRE(s.data && s.n > 0 && s.c > 0, "factor job %d segment %d: empty", ji, si);
      DevSegment ds;
      memset(&ds, 0, sizeof(ds));
      KFB_REQUIRE(s.dtype >= KFB_F32 && s.dtype <= KFB_BF16, "unknown segment dtype %d", s.dtype);
      ds.data = s.data; ds.dtype = s.dtype; ds.kind = s.kind; ds.has_bias = s.has_bias; ds.c = (int)s.c;
      ds.h = (int)s.h; ds.w = (int)s.w; ds.ld = s.ld; ds.weight = s.weight; ds.dev_scale = s.dev_scale;
      ds.kh = ds.kw = ds.sh = ds.sw = 1;
      int feat;
      switch (s.kind) {
        case KFB_SEG_ROWS:
          KFB_REQUIRE(s.ld >= s.c, "segment pitch %lld < feature count %lld", (long long)s.ld, (long long)s.c);
          ds.K = s.n; ds.oh = 1; ds.ow = 1; feat = (int)s.c + (s.has_bias ? 1 : 0);
          break;
        case KFB_SEG_NCHW:
          KFB_REQUIRE(!s.has_bias && s.h > 0 && s.w > 0, "NCHW segment cannot carry a bias column");
          ds.oh = (int)s.h; ds.ow = (int)s.w; ds.K = s.n * s.h * s.w; feat = (int)s.c;
          break;
        case KFB_SEG_IM2COL:
          KFB_REQUIRE(s.kh > 0 && s.kw > 0 && s.sh > 0 && s.sw > 0 && s.ph >= 0 && s.pw >= 0, "bad conv geometry");
          ds.kh = s.kh; ds.kw = s.kw; ds.sh = s.sh; ds.sw = s.sw; ds.ph = s.ph; ds.pw = s.pw;
          ds.oh = (int)((s.h + 2 * s.ph - s.kh) / s.sh + 1);
          ds.ow = (int)((s.w + 2 * s.pw - s.kw) / s.sw + 1);
          KFB_REQUIRE(ds.oh > 0 && ds.ow > 0, "conv output is empty");
          KFB_REQUIRE(s.n * s.c * s.h * s.w < (int64_t)1 << 40, "tensor too large");
          ds.K = s.n * ds.oh * ds.ow; feat = (int)s.c * s.kh * s.kw + (s.has_bias ? 1 : 0);
          break;
        default:
          KFB_FAIL(KFB_ERR_INVALID, "unknown segment kind %d", s.kind);
      }
      KFB_REQUIRE(feat == j.d, "factor job %d segment %d: feature count %d != d %d", ji, si, feat, j.d);
      k_total += ds.K;
      KFB_REQUIRE(ds.K < ((int64_t)1 << 31), "segment has %lld rows; at most 2^31-1 are supported", (long long)ds.K);
      // row-major captures (Linear layers) reach the tensor-core SYRK through a transposed fp32 copy (PadJob
      // variant 5) that TMA reads like a one-image NCHW capture
      if (s.kind == KFB_SEG_ROWS &&
          !(stage_mode() && tma_mode() && rows_tc_mode() && s.ld < ((int64_t)1 << 31) &&
            (int64_t)feat * ((s.n + 31) / 32 * 32) < ((int64_t)1 << 31)))
        tc = false;
      if (s.c * s.h * s.w >= ((int64_t)1 << 26)) tc = false;   // packed row offsets hold 26 bits
      ds.tma = -1;
      ds.nb = 0;
      ds.nomask = 0;
      magic_div((int64_t)ds.oh * ds.ow, &ds.mul_s, &ds.sh_s);
      magic_div(ds.ow, &ds.mul_w, &ds.sh_w);
      plan->segs.push_back(ds);
    }
    // split K so that the job alone offers ~4 CTAs per SM, chunks of at least 256 rows
    int want = (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(k_total, 256), ceil_div(target, dj.n_pairs)));
    int chunks = 0;
    for (int si = 0; si < j.n_segments; ++si) {
      DevSegment& ds = plan->segs[dj.seg_start + si];
      int64_t n = std::max<int64_t>(1, std::min<int64_t>(ceil_div64(ds.K, 256), ceil_div64((int64_t)want * ds.K, k_total)));
      ds.chunk_len = ceil_div64(ceil_div64(ds.K, n), BK) * BK;
      ds.n_chunks = (int)ceil_div64(ds.K, ds.chunk_len);
      ds.chunk_start = chunks;
      chunks += ds.n_chunks;
    }
    dj.n_chunks = chunks;
    dj.use_tc = tc ? 1 : 0;
    // 16-bit operands for kind::f16 need the TMA-fed single-CTA kernel
    const bool half16 = tc && half_mode() && tma_mode() && stage_mode() &&
                        !(syrk_pair_mode() > 0 && ceil_div(j.d, 128) >= syrk_pair_mode());
    for (int si = 0; si < j.n_segments; ++si) {
      DevSegment& ds = plan->segs[dj.seg_start + si];
      plan->seg_pad.push_back(-1);
      if (!tc) continue;
      bool aligned = ((uintptr_t)ds.data & 15) == 0;
      if (ds.kind == KFB_SEG_IM2COL && stage_mode()) {
        // read this capture from an fp32 staging copy instead (see PadJob)
        PadJob pj;
        memset(&pj, 0, sizeof(pj));
        pj.src = ds.data; pj.dtype = ds.dtype; pj.h = ds.h; pj.w = ds.w;
        pj.pad_h = ds.ph; pj.pad_w = ds.pw; pj.ow_valid = ds.ow;
        const int64_t images = ds.K / ((int64_t)ds.oh * ds.ow);
        const bool pointwise = ds.kh == 1 && ds.kw == 1;
        bool keep16 = false;
        if (pointwise) {
          pj.kw = 1; pj.sy = ds.sh; pj.sx = ds.sw;
          pj.hp = ds.oh; pj.wp = ds.ow;
        } else {
          pj.kw = ds.kw; pj.sy = 1; pj.sx = ds.sw;
          pj.hp = (ds.oh - 1) * ds.sh + ds.kh;
          // a 16-bit capture keeps its type in the planes (eight pixels per 16-byte group)
          keep16 = half16 && ds.dtype != KFB_F32;
          pj.wp = keep16 ? (ds.ow + 7) / 8 * 8 : (ds.ow + 3) / 4 * 4;
          pj.out16 = keep16 ? 1 : 0;
        }
        pj.total = images * ds.c * pj.kw * pj.hp * pj.wp;
        magic_div(pj.wp, &pj.mul_wp, &pj.sh_wp);
        magic_div(pj.hp, &pj.mul_hp, &pj.sh_hp);
        magic_div(pj.kw, &pj.mul_kw, &pj.sh_kw);
        magic_div(pj.wp / 4, &pj.mul_wq, &pj.sh_wq);
        pj.variant = 0;
        if (!pointwise && pj.sy == 1) {
          if (pj.sx == 1 && pj.kw == 3) pj.variant = 1;
          else if (pj.sx == 2 && pj.kw == 3) pj.variant = 2;
          else if (pj.sx == 2 && pj.kw == 7) pj.variant = 3;
        }
        pj.quads = images * ds.c * pj.hp * (pj.wp / 4);
        if (pj.total < ((int64_t)1 << 31) && (int64_t)ds.c * pj.kw * pj.hp * pj.wp < ((int64_t)1 << 26) &&
            images * ds.oh * pj.wp < ((int64_t)1 << 31)) {
          plan->seg_pad.back() = (int)plan->pads.size();
          plan->pad_off.push_back(plan->stage_floats);
          plan->stage_floats += align_up((size_t)(keep16 ? (pj.total + 1) / 2 : pj.total), 64);
          plan->pads.push_back(pj);
          ds.fmt = keep16 ? ds.dtype : 0;
          ds.dtype = KFB_F32;
          ds.h = pj.hp; ds.w = keep16 ? pj.wp / 2 : pj.wp;      // 16-bit planes: everything in 4-byte words (fmt)
          ds.ph = ds.pw = 0;
          if (pointwise) {
            ds.sh = ds.sw = 1;
            if (!ds.has_bias) ds.kind = KFB_SEG_NCHW;
          } else {
            ds.nomask = 1;
            ds.ow_valid = ds.ow;                 // in PIXELS also for 16-bit planes (two per word)
            ds.ow = ds.w;                        // X rows now count padded pixels (pixel pairs)
            ds.K = images * ds.oh * ds.ow;
            ds.sw = 1;                           // the x-stride is applied by the copy
          }
          magic_div((int64_t)ds.oh * ds.ow, &ds.mul_s, &ds.sh_s);
          magic_div(ds.ow, &ds.mul_w, &ds.sh_w);
          aligned = true;       // staging copies start on 256-byte boundaries
        }
      }
      if (ds.kind == KFB_SEG_ROWS) {
        // X (rows x features, any pitch, any type) -> fp32 [features (+ a row of ones for the bias)][Kp], Kp = rows
        // rounded up to 32 and zero-filled: from here on a plain NCHW segment of one "image" of Kp "pixels"
        // A 16-bit capture keeps its type (DevSegment::fmt): Kp is then a multiple of 64 rows = 32 words.
        const bool keep16 = half16 && ds.dtype != KFB_F32;
        const int64_t Kp = keep16 ? (ds.K + 63) / 64 * 64 : (ds.K + 31) / 32 * 32;
        const int C = ds.c + (ds.has_bias ? 1 : 0);
        PadJob pj;
        memset(&pj, 0, sizeof(pj));
        pj.src = ds.data; pj.dtype = ds.dtype; pj.variant = 5; pj.out16 = keep16 ? 1 : 0;
        pj.channels = C; pj.h = ds.c; pj.w = (int)ds.ld; pj.hp = (int)ds.K; pj.wp = (int)Kp; pj.kw = 1;
        pj.quads = Kp / 32;                                   // 32 x 32 tiles along K
        pj.total = (int64_t)C * Kp;
        plan->seg_pad.back() = (int)plan->pads.size();
        plan->pad_off.push_back(plan->stage_floats);
        plan->stage_floats += align_up((size_t)(keep16 ? (pj.total + 1) / 2 : pj.total), 64);
        plan->pads.push_back(pj);
        const int64_t Kw = keep16 ? Kp / 2 : Kp;              // row length in 4-byte words
        ds.fmt = keep16 ? ds.dtype : 0;
        ds.kind = KFB_SEG_NCHW; ds.dtype = KFB_F32; ds.has_bias = 0;
        ds.c = C; ds.h = 1; ds.w = (int)Kw; ds.oh = 1; ds.ow = (int)Kw; ds.K = Kw; ds.ld = 0;
        magic_div(ds.K, &ds.mul_s, &ds.sh_s);
        magic_div(ds.ow, &ds.mul_w, &ds.sh_w);
        aligned = true;
      }
      if (half16 && ds.kind == KFB_SEG_NCHW && ds.dtype != KFB_F32 && plan->seg_pad.back() < 0) {
        // 16-bit NCHW captures (AMP) feed kind::f16 MMAs.  The operand is described to TMA and to the kernel as an
        // fp32 matrix of HALF as many pixels (DevSegment::fmt): image rows that are 16-byte multiples (S % 8 == 0)
        // are read in place; the others (14 x 14, 7 x 7 maps) through a 16-bit [channels][images * wp] copy.
        const int64_t S0 = (int64_t)ds.oh * ds.ow, images = ds.K / S0, wp = (S0 + 7) / 8 * 8;
        if ((S0 & 7) == 0 && S0 >= 2 * BK && aligned) {
          ds.fmt = ds.dtype; ds.dtype = KFB_F32;
          ds.h = 1; ds.w = (int)(S0 / 2); ds.oh = 1; ds.ow = ds.w;
          ds.K = images * (S0 / 2);
          magic_div((int64_t)ds.oh * ds.ow, &ds.mul_s, &ds.sh_s);
          magic_div(ds.ow, &ds.mul_w, &ds.sh_w);
        } else if (images * wp >= 2 * BK && (int64_t)ds.c * images * wp < ((int64_t)1 << 31)) {
          PadJob pj;
          memset(&pj, 0, sizeof(pj));
          pj.src = ds.data; pj.dtype = ds.dtype; pj.variant = 4; pj.out16 = 1;
          pj.channels = ds.c; pj.hp = (int)images; pj.wp = (int)wp; pj.ow_valid = (int)S0; pj.kw = 1;
          pj.total = (int64_t)ds.c * images * wp;
          magic_div(pj.wp, &pj.mul_wp, &pj.sh_wp);
          magic_div(pj.hp, &pj.mul_hp, &pj.sh_hp);
          plan->seg_pad.back() = (int)plan->pads.size();
          plan->pad_off.push_back(plan->stage_floats);
          plan->stage_floats += align_up((size_t)((pj.total + 1) / 2), 64);
          plan->pads.push_back(pj);
          ds.fmt = ds.dtype; ds.dtype = KFB_F32;
          ds.h = 1; ds.w = (int)(images * wp / 2);
          ds.oh = 1; ds.ow = ds.w;
          ds.K = ds.w;
          magic_div(ds.K, &ds.mul_s, &ds.sh_s);
          magic_div(ds.ow, &ds.mul_w, &ds.sh_w);
          aligned = true;
        }
      }
      {
        // NCHW operands TMA cannot take as they are: stage an fp32 [channels][images * padded pixels]
        // copy (PadJob variant 4), which TMA reads as one "image" of K' = images * wp pixels
        const int64_t S0 = (int64_t)ds.oh * ds.ow;
        const bool tma_ok = ds.dtype == KFB_F32 && (S0 & 3) == 0 && S0 >= BK && aligned;
        const int64_t images = ds.K / S0, wp = (S0 + 3) / 4 * 4;
        if (ds.kind == KFB_SEG_NCHW && !tma_ok && stage_mode() && tma_mode() && plan->seg_pad.back() < 0 &&
            images * wp >= BK && (int64_t)ds.c * images * wp < ((int64_t)1 << 31)) {
          PadJob pj;
          memset(&pj, 0, sizeof(pj));
          pj.src = ds.data; pj.dtype = ds.dtype; pj.variant = 4;
          pj.channels = ds.c; pj.hp = (int)images; pj.wp = (int)wp; pj.ow_valid = (int)S0; pj.kw = 1;
          pj.total = (int64_t)ds.c * images * wp;
          magic_div(pj.wp, &pj.mul_wp, &pj.sh_wp);
          magic_div(pj.hp, &pj.mul_hp, &pj.sh_hp);
          plan->seg_pad.back() = (int)plan->pads.size();
          plan->pad_off.push_back(plan->stage_floats);
          plan->stage_floats += align_up((size_t)pj.total, 64);
          plan->pads.push_back(pj);
          ds.dtype = KFB_F32;
          ds.h = 1; ds.w = (int)(images * wp);
          ds.oh = 1; ds.ow = ds.w;
          ds.K = images * wp;
          magic_div(ds.K, &ds.mul_s, &ds.sh_s);
          magic_div(ds.ow, &ds.mul_w, &ds.sh_w);
          aligned = true;
        }
      }
      // fp32 NCHW operands whose image rows are 16-byte multiples are fetched by TMA (3-D map
      // (pixel, channel, image), box 32 x 128 x 1); everything else is gathered by the producer warps
      const int64_t S = (int64_t)ds.oh * ds.ow;
      if (tma_mode() && ds.kind == KFB_SEG_NCHW && ds.dtype == KFB_F32 && (S & 3) == 0 && S >= BK && aligned) {
        ds.nb = (int)ceil_div64(S, BK);
        ds.tma = (int)plan->tmaps.size();
        plan->tmaps.emplace_back();
        plan->tmap_ptr.push_back(nullptr);
      }
    }
    // short-K single-segment jobs (the 7x7-map layers: the LARGEST factors) are one work item per
    // block: their SYRK epilogue fuses scale + running average + mirror + triangle export
    dj.fused = (tc && j.n_segments == 1 && k_total <= fused_max_rows()) ? 1 : 0;
    // a fused job's packed triangle is exported from its finished state by the (coalesced) pack kernel
    // after the SYRK rather than row by row from the epilogue's registers
    if (dj.fused && j.write_state) dj.tri_out = nullptr;
    dj.T128 = ceil_div(j.d, 128);
    dj.native = (tc && !dj.fused) ? 1 : 0;
    // split-K jobs of the single-CTA tensor-core kernel can finish themselves (DevJob::la); their accumulators
    // and counters form the second part of the scratch region
    const int pair_min = syrk_pair_mode();
    dj.la = (dj.native && last_arriver_mode() && !stream_k_mode() && !(pair_min > 0 && dj.T128 >= pair_min)) ? 1 : 0;
    dj.la_items = 0; dj.arrive = nullptr; dj.arrive_off = 0;
    if (dj.la) {
      const int TB = (dj.T128 + 1) / 2;
      dj.scratch_off = la_floats;
      la_floats += (size_t)dj.T128 * dj.T128 * 128 * 128;
      dj.arrive_off = la_floats;
      la_floats += align_up((size_t)TB * TB, 64);
      plan->la_blocks += TB * (TB + 1) / 2;
      if (j.write_state) dj.tri_out = nullptr;    // exported from the finished state by the pack kernel, like a fused job's
    } else {
      dj.scratch_off = plan->scratch_floats;
      if (dj.native) plan->scratch_floats += (size_t)dj.T128 * dj.T128 * 128 * 128;
      else if (!dj.fused) plan->scratch_floats += align_up((size_t)dj.ld_scratch * j.d, 64);
    }
    plan->item_start.push_back(items);
    plan->fin_start.push_back(fins);
    if (!tc) items += dj.n_pairs * chunks;
    if (!dj.fused && !dj.la) fins += dj.T32 * (dj.T32 + 1) / 2;
    plan->jobs.push_back(dj);
  }
  plan->la_begin = plan->scratch_floats;
  for (DevJob& dj : plan->jobs)
    if (dj.la) { dj.scratch_off += plan->la_begin; dj.arrive_off += plan->la_begin; }
  // the queue of completed blocks: [pushed, popped, CTAs gone, ...][la_blocks slots]
  plan->la_ctl_off = plan->la_begin + la_floats;
  plan->la_ctl = allow_tc && last_arriver_mode() && !stream_k_mode();
  if (plan->la_ctl) la_floats += align_up((size_t)kLaCtlInts + plan->la_blocks, 64);
  plan->scratch_floats += la_floats;
  static uint64_t generation = 0;
  plan->generation = ++generation;
  plan->item_start.push_back(items);
  plan->fin_start.push_back(fins);
  plan->n_items = items;
  plan->n_fin = fins;
  // workspace layout: [jobs][segments][item_start][fin_start][scratch]
  size_t off = 0;
  plan->off_jobs = off;   off = align_up(off + plan->jobs.size() * sizeof(DevJob), 256);
  plan->off_segs = off;   off = align_up(off + plan->segs.size() * sizeof(DevSegment), 256);
  plan->off_items = off;  off = align_up(off + plan->item_start.size() * sizeof(int), 256);
  plan->off_fin = off;    off = align_up(off + plan->fin_start.size() * sizeof(int), 256);
  plan_syrk_tc(plan, num_sms);
  plan->off_tc_items = off; off = align_up(off + plan->tc_items.size() * sizeof(TcItem), 256);
  plan->off_cta_first = off; off = align_up(off + plan->cta_first.size() * sizeof(int), 256);
  plan->off_tmaps = off;    off = align_up(off + plan->tmaps.size() * sizeof(CUtensorMap), 256);
  // staging copies: blocks of 256 threads x 4 floats
  int pad_blocks = 0;
  for (const PadJob& pj : plan->pads) {
    plan->pad_start.push_back(pad_blocks);
    pad_blocks += (int)(pj.variant == 5 ? pj.quads * ceil_div(pj.channels, 32)
                        : (pj.variant >= 1 && pj.variant <= 3) ? ceil_div64(pj.quads, 256) : ceil_div64(pj.total, 1024));
  }
  plan->pad_start.push_back(pad_blocks);
  plan->off_pads = off;      off = align_up(off + plan->pads.size() * sizeof(PadJob), 256);
  plan->off_pad_start = off; off = align_up(off + plan->pad_start.size() * sizeof(int), 256);
  plan->off_scratch = off;   off = align_up(off + plan->scratch_floats * sizeof(float), 256);
  plan->off_stage = off;
  plan->total_bytes = off + plan->stage_floats * sizeof(float);
  return KFB_OK;
}

// ---- plan cache ---------------------------------------------------------------------------------
// Planning (tile/chunk tables, the sorted tensor-core work list) depends only on the GEOMETRY of the
// jobs, which is the same at every factor update of a training run; only pointers, weights and the
// seed/write flags change.  The last plan is kept and patched instead of being rebuilt twice per
// update (once for the workspace size, once for the launch).
namespace {
std::mutex g_plan_mu;
FactorPlan g_plan;
std::vector<int64_t> g_plan_sig;

void signature(const kfb_factor_job* jobs, int n_jobs, int num_sms, bool tc, std::vector<int64_t>* sig) {
  sig->clear();
  sig->push_back(n_jobs); sig->push_back(num_sms); sig->push_back(tc); sig->push_back(tma_mode()); sig->push_back(stage_mode()); sig->push_back(syrk_pair_mode()); sig->push_back(fused_max_rows()); sig->push_back(stream_k_mode()); sig->push_back(last_arriver_mode()); sig->push_back(rows_tc_mode()); sig->push_back(half_mode());
  for (int i = 0; i < n_jobs; ++i) {
    const kfb_factor_job& j = jobs[i];
    sig->push_back(j.d); sig->push_back(j.n_segments); sig->push_back(j.ld_state);
    if (!j.segments) continue;
    for (int k = 0; k < j.n_segments; ++k) {
      const kfb_segment& s = j.segments[k];
      const int64_t v[] = {s.kind, s.has_bias, s.n, s.c, s.h, s.w, s.kh, s.kw, s.sh, s.sw, s.ph, s.pw, s.ld, s.dtype,
                           (int64_t)((uintptr_t)s.data & 15)};   // alignment decides TMA eligibility
      sig->insert(sig->end(), v, v + sizeof(v) / sizeof(v[0]));
    }
  }
}
}  // namespace

// Returns the (cached or freshly built) plan with this call's pointers patched in.  The caller
// must hold g_plan_mu for as long as it uses the plan.
static int cached_plan(const kfb_factor_job* jobs, int n_jobs, int num_sms, bool tc, FactorPlan** out) {
  KFB_REQUIRE(n_jobs >= 0 && (jobs || n_jobs == 0), "bad job array");
  std::vector<int64_t> sig;
  signature(jobs, n_jobs, num_sms, tc, &sig);
  if (sig != g_plan_sig) {
    g_plan_sig.clear();
    int rc = plan_factor(jobs, n_jobs, num_sms, tc, &g_plan);
    if (rc) return rc;
    g_plan_sig.swap(sig);
  } else {
    for (int i = 0; i < n_jobs; ++i) {
      const kfb_factor_job& j = jobs[i];
      KFB_REQUIRE(j.state, "factor job %d: null state", i);
      DevJob& dj = g_plan.jobs[i];
      dj.state = j.state; dj.seed_identity = j.seed_identity; dj.write_state = j.write_state;
      dj.dev_rescale = j.dev_rescale;
      dj.tri_out = ((dj.fused || dj.la) && j.write_state) ? nullptr : j.tri_out;
      for (int k = 0; k < j.n_segments; ++k) {
        KFB_REQUIRE(j.segments[k].data, "factor job %d segment %d: null data", i, k);
        DevSegment& ds = g_plan.segs[dj.seg_start + k];
        const int pad = g_plan.seg_pad[dj.seg_start + k];
        if (pad >= 0) g_plan.pads[pad].src = j.segments[k].data;   // ds.data is the staging copy
        else ds.data = j.segments[k].data;
        ds.weight = j.segments[k].weight;
        ds.dev_scale = j.segments[k].dev_scale;
      }
    }
  }
  *out = &g_plan;
  return KFB_OK;
}

// Workspaces the caller has promised to use for kfb_factor_update only (kfb_factor_workspace_dedicate): what the
// library leaves in them survives until its next call.  Guarded by g_plan_mu.
namespace {
struct Dedicated {
  void* ptr;
  size_t bytes;
  uint64_t clean_for;     // generation of the plan whose last-arriver region is known to be zero; 0 = unknown
};
std::vector<Dedicated> g_dedicated;
Dedicated* find_dedicated(void* p) {
  for (Dedicated& d : g_dedicated)
    if (d.ptr == p) return &d;
  return nullptr;
}
}  // namespace

int dedicate_workspace(void* workspace, size_t bytes, bool on) {
  std::lock_guard<std::mutex> lock(g_plan_mu);
  for (size_t i = 0; i < g_dedicated.size(); ++i)
    if (g_dedicated[i].ptr == workspace) { g_dedicated.erase(g_dedicated.begin() + i); break; }
  if (on) g_dedicated.push_back(Dedicated{workspace, bytes, 0});
  return KFB_OK;
}

int factor_update_with(const kfb_factor_job* jobs, int n_jobs, float alpha, Context* ctx, void* workspace,
                       size_t workspace_bytes, const char* what, cudaStream_t stream) {
  std::lock_guard<std::mutex> lock(g_plan_mu);
  FactorPlan* plan;
  int rc = cached_plan(jobs, n_jobs, ctx->num_sms, ctx->umma_ok && mma_mode() != 0, &plan);
  if (rc) return rc;
  if (!workspace || workspace_bytes < plan->total_bytes)
    KFB_FAIL(KFB_ERR_WORKSPACE, "%s: these %d job(s) need %zu bytes of scratch but %zu are available", what, n_jobs,
             plan->total_bytes, workspace_bytes);
  KFB_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  bind_workspace(plan, workspace);
  rc = refresh_tmaps(plan, ctx);
  if (rc) return rc;
  // a dedicated workspace that last served this very plan holds zeroed last-arriver accumulators and counters
  Dedicated* own = find_dedicated(workspace);
  const bool la_clean = own && own->clean_for == plan->generation && own->bytes >= plan->total_bytes;
  if (own) own->clean_for = 0;
  rc = upload_plan(plan, workspace, stream, la_clean);
  if (rc) return rc;
  rc = launch_pad(*plan, workspace, stream);
  if (rc) return rc;
  rc = launch_syrk(*plan, workspace, ctx, alpha, stream);
  if (rc) return rc;
  if (own) own->clean_for = plan->generation;
  rc = launch_finalize(*plan, workspace, alpha, stream);
  if (rc) return rc;
  std::vector<kfb_tri_job> deferred;
  for (int i = 0; i < n_jobs; ++i)
    if ((plan->jobs[i].fused || plan->jobs[i].la) && jobs[i].write_state && jobs[i].tri_out) {
      kfb_tri_job t;
      t.tri = jobs[i].tri_out; t.state = jobs[i].state; t.ld_state = jobs[i].ld_state; t.d = jobs[i].d;
      deferred.push_back(t);
    }
  if (!deferred.empty()) return run_tri(deferred.data(), (int)deferred.size(), 1.f, false, stream);
  return KFB_OK;
}

size_t factor_workspace_bytes(const kfb_factor_job* jobs, int n_jobs, Context* ctx) {
  std::lock_guard<std::mutex> lock(g_plan_mu);
  FactorPlan* plan;
  bool tc = ctx && ctx->umma_ok && mma_mode() != 0;
  if (cached_plan(jobs, n_jobs, ctx ? ctx->num_sms : 148, tc, &plan) != KFB_OK) return 0;
  return plan->total_bytes;
}

// Point the plan's scratch accumulators and staging copies into this call's workspace.
void bind_workspace(FactorPlan* plan, void* workspace) {
  char* ws = (char*)workspace;
  float* scratch = (float*)(ws + plan->off_scratch);
  for (auto& j : plan->jobs) {
    j.scratch = scratch + j.scratch_off;
    j.arrive = j.la ? (int*)(scratch + j.arrive_off) : nullptr;
  }
  float* stage = (float*)(ws + plan->off_stage);
  for (size_t i = 0; i < plan->segs.size(); ++i) {
    const int pad = plan->seg_pad[i];
    if (pad < 0) continue;
    plan->pads[pad].dst = stage + plan->pad_off[pad];
    plan->segs[i].data = plan->pads[pad].dst;
  }
}

int upload_plan(FactorPlan* plan, void* workspace, cudaStream_t stream, bool la_clean) {
  char* ws = (char*)workspace;
  float* scratch = (float*)(ws + plan->off_scratch);
  if (!plan->pads.empty()) {
    KFB_CUDA(cudaMemcpyAsync(ws + plan->off_pads, plan->pads.data(), plan->pads.size() * sizeof(PadJob), cudaMemcpyHostToDevice, stream));
    KFB_CUDA(cudaMemcpyAsync(ws + plan->off_pad_start, plan->pad_start.data(), plan->pad_start.size() * sizeof(int), cudaMemcpyHostToDevice, stream));
  }
  KFB_CUDA(cudaMemcpyAsync(ws + plan->off_jobs, plan->jobs.data(), plan->jobs.size() * sizeof(DevJob), cudaMemcpyHostToDevice, stream));
  KFB_CUDA(cudaMemcpyAsync(ws + plan->off_segs, plan->segs.data(), plan->segs.size() * sizeof(DevSegment), cudaMemcpyHostToDevice, stream));
  KFB_CUDA(cudaMemcpyAsync(ws + plan->off_items, plan->item_start.data(), plan->item_start.size() * sizeof(int), cudaMemcpyHostToDevice, stream));
  KFB_CUDA(cudaMemcpyAsync(ws + plan->off_fin, plan->fin_start.data(), plan->fin_start.size() * sizeof(int), cudaMemcpyHostToDevice, stream));
  if (!plan->tc_items.empty())
    KFB_CUDA(cudaMemcpyAsync(ws + plan->off_tc_items, plan->tc_items.data(), plan->tc_items.size() * sizeof(TcItem), cudaMemcpyHostToDevice, stream));
  if (!plan->cta_first.empty())
    KFB_CUDA(cudaMemcpyAsync(ws + plan->off_cta_first, plan->cta_first.data(), plan->cta_first.size() * sizeof(int), cudaMemcpyHostToDevice, stream));
  if (!plan->tmaps.empty())
    KFB_CUDA(cudaMemcpyAsync(ws + plan->off_tmaps, plan->tmaps.data(), plan->tmaps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice, stream));
  // the last-arriver part is re-zeroed by the kernel that used it: cleared here only when this workspace is not
  // known to hold what the previous update of the same plan left (kfb_factor_workspace_dedicate)
  const size_t clear = la_clean ? plan->la_begin : plan->scratch_floats;
  if (clear) KFB_CUDA(cudaMemsetAsync(scratch, 0, clear * sizeof(float), stream));
  return KFB_OK;
}

int launch_syrk_simt(const FactorPlan& plan, void* workspace, cudaStream_t stream) {
  if (plan.n_items == 0) return KFB_OK;
  char* ws = (char*)workspace;
  syrk_simt_kernel<<<plan.n_items, NT, 0, stream>>>((const DevJob*)(ws + plan.off_jobs),
                                                   (const DevSegment*)(ws + plan.off_segs),
                                                   (const int*)(ws + plan.off_items),
                                                   (int)plan.jobs.size());
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

int launch_pad(const FactorPlan& plan, void* workspace, cudaStream_t stream) {
  if (plan.pads.empty() || plan.pad_start.back() == 0) return KFB_OK;
  char* ws = (char*)workspace;
  pad_kernel<<<plan.pad_start.back(), 256, 0, stream>>>((const PadJob*)(ws + plan.off_pads),
                                                       (const int*)(ws + plan.off_pad_start), (int)plan.pads.size());
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

int launch_finalize(const FactorPlan& plan, void* workspace, float alpha, cudaStream_t stream) {
  if (plan.n_fin == 0) return KFB_OK;
  char* ws = (char*)workspace;
  finalize_kernel<<<plan.n_fin, 256, 0, stream>>>((const DevJob*)(ws + plan.off_jobs),
                                                 (const int*)(ws + plan.off_fin),
                                                 (int)plan.jobs.size(), alpha);
  KFB_CHECK_LAUNCH();
  return KFB_OK;
}

}  // namespace kfb

// ============================================================================================
// C ABI
// ============================================================================================
extern "C" {

size_t kfb_factor_workspace_bytes(const kfb_factor_job* jobs, int32_t n_jobs) {
  return kfb::factor_workspace_bytes(jobs, n_jobs, kfb::current_context());
}

int kfb_factor_update(const kfb_factor_job* jobs, int32_t n_jobs, float alpha, void* workspace,
                      size_t workspace_bytes, void* stream_) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  if (n_jobs == 0) return KFB_OK;
  KFB_REQUIRE(alpha >= 0.f && alpha <= 1.f, "alpha must be in [0, 1], got %g", alpha);   // 0: state <- F
  if (alpha == 1.f) return KFB_OK;   // update_running_avg is a no-op (kfac/layers/utils.py:175)
  return kfb::factor_update_with(jobs, n_jobs, alpha, ctx, workspace, workspace_bytes, "kfb_factor_update",
                                 (cudaStream_t)stream_);
}

int kfb_factor_workspace_dedicate(void* workspace, size_t workspace_bytes) {
  KFB_REQUIRE(workspace && ((uintptr_t)workspace & 255) == 0, "workspace must be a 256-byte aligned device pointer");
  return kfb::dedicate_workspace(workspace, workspace_bytes, true);
}

int kfb_factor_workspace_release(void* workspace) { return kfb::dedicate_workspace(workspace, 0, false); }

int kfb_tri_unpack(const kfb_tri_job* jobs, int32_t n_jobs, float scale, void* stream) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  return kfb::run_tri(jobs, n_jobs, scale, true, (cudaStream_t)stream);
}

int kfb_tri_pack(const kfb_tri_job* jobs, int32_t n_jobs, void* stream) {
  kfb::Context* ctx;
  int rc = kfb::require_context(&ctx);
  if (rc) return rc;
  return kfb::run_tri(jobs, n_jobs, 1.f, false, (cudaStream_t)stream);
}

}  // extern "C"
