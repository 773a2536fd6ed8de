// Shared host/device helpers for libkfac_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/kfac_b200.h"

#define KFB_HIDDEN __attribute__((visibility("hidden")))

namespace kfb {

// ---- error reporting -----------------------------------------------------------------
KFB_HIDDEN void set_error(const char* fmt, ...);
KFB_HIDDEN const char* get_error();
KFB_HIDDEN void count_launch();

#define KFB_FAIL(code, ...)           \
  do {                                \
    ::kfb::set_error(__VA_ARGS__);    \
    return (code);                    \
  } while (0)

#define KFB_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      ::kfb::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,              \
                       cudaGetErrorString(_e));                                          \
      return KFB_ERR_CUDA;                                                               \
    }                                                                                    \
  } while (0)

// every kernel of this library is launched through this macro, so the counter is the number
// of OUR kernels launched (cuSOLVER's own kernels and memcpy/memset nodes are not counted)
#define KFB_CHECK_LAUNCH()            \
  do {                                \
    ::kfb::count_launch();            \
    KFB_CUDA(cudaGetLastError());     \
  } while (0)

#define KFB_REQUIRE(cond, ...)                      \
  do {                                              \
    if (!(cond)) KFB_FAIL(KFB_ERR_INVALID, __VA_ARGS__); \
  } while (0)

// ---- per-device context ----------------------------------------------------------------
struct Context {
  int device = -1;
  int num_sms = 0;
  int cc_major = 0, cc_minor = 0;
  void* cusolver = nullptr;       // cusolverDnHandle_t
  void* encode_tiled = nullptr;   // cuTensorMapEncodeTiled entry point (driver API, via cudart)
  bool umma_ok = false;           // sm_100 device and TMA entry point available
};

KFB_HIDDEN Context* current_context();          // context selected by the last kfb_init on this thread
KFB_HIDDEN int require_context(Context** out);  // KFB_OK or error if kfb_init was never called
KFB_HIDDEN int mma_mode();                      // 0 SIMT, 1 tcgen05 (TF32), 2 tcgen05 with a 3xTF32 precondition chain
KFB_HIDDEN int syrk_pair_mode();                // factor SYRK: factors of at least this many 128-wide tiles per side go to the CTA-pair (cta_group::2) kernel; 0 (default): none.  env KFB_SYRK_2CTA=<tiles>
KFB_HIDDEN int stage_mode();                    // factor SYRK: 1 (default) im2col operands are read from zero-padded fp32 staging copies; env KFB_SYRK_STAGE=0: in place
KFB_HIDDEN void set_stage_override(int v);       // this thread: -1 follow the environment, 0 / 1 force staging off / on
KFB_HIDDEN int tma_mode();                      // factor SYRK: 1 (default) feeds eligible NCHW operands by TMA; env KFB_SYRK_TMA=0: gathers only

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// ---- batched launches ---------------------------------------------------------------------
// Job descriptors travel as a by-value kernel parameter (CUDA >= 12.1 allows ~32 KB), so a
// whole stage over all layers is ONE launch with no descriptor upload.  `start[i]` is the
// first block of job i; blocks find their job with a binary search.
template <typename Job, int kMaxJobs>
struct Batch {
  int n;
  int start[kMaxJobs + 1];
  Job jobs[kMaxJobs];
};

template <typename BatchT>
__device__ __forceinline__ int find_job(const BatchT& b, int block) {
  int lo = 0, hi = b.n;   // invariant: start[lo] <= block < start[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (b.start[mid] <= block) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace kfb
