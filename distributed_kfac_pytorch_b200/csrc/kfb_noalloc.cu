// kfb_factor_update_static: the factor update for callers that have no device memory to lend.
//
// Some hosts of this library cannot give it a workspace: an embedding runtime that carved up
// the whole GPU at start-up and hands plugins no allocator, or a service whose memory budget
// is fixed once it is serving and where a cudaMalloc (which may stall the device or fail) is not
// acceptable in the step loop.  For them the split-K accumulators and job descriptors that
// kfb_factor_update keeps in caller scratch live here instead, in a fixed array that is part of
// the library image.  The price is that every GPU that loads libkfac_b200.so gives up
// KFB_STATIC_SCRATCH_BYTES of memory until the process exits, used or not.
#include "kfb_factor.cuh"

#include <mutex>

namespace kfb {
namespace {

// Device-side storage.  In host code this name is only a handle for the runtime's symbol
// table - it is NOT a device address and must never be passed to a kernel; the per-device
// address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_static_scratch[KFB_STATIC_SCRATCH_BYTES];

static void* g_scratch_ptr[64];
static std::mutex g_scratch_mu;

// Resolve the array's address on the CURRENT device, once per device.
int static_scratch(int device, void** out) {
  KFB_REQUIRE(device >= 0 && device < 64, "device index %d out of range", device);
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (!g_scratch_ptr[device]) {
    void* p = nullptr;
    KFB_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    KFB_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
    g_scratch_ptr[device] = p;
  }
  *out = g_scratch_ptr[device];
  return KFB_OK;
}

}  // namespace
}  // namespace kfb

extern "C" {

size_t kfb_static_scratch_bytes(void) { return KFB_STATIC_SCRATCH_BYTES; }

int kfb_factor_update_static(const kfb_factor_job* jobs, int32_t n_jobs, float alpha, void* stream_) {
  using namespace kfb;
  Context* ctx;
  int rc = require_context(&ctx);
  if (rc) return rc;
  if (n_jobs == 0) return KFB_OK;
  KFB_REQUIRE(alpha > 0.f && alpha <= 1.f, "alpha must be in (0, 1], got %g", alpha);
  if (alpha == 1.f) return KFB_OK;
  int device = -1;
  KFB_CUDA(cudaGetDevice(&device));
  KFB_REQUIRE(device == ctx->device, "current device %d is not the device of the kfb context (%d)", device,
              ctx->device);
  size_t need = factor_workspace_bytes(jobs, n_jobs, ctx);
  if (need == 0) return KFB_ERR_INVALID;      // the planner has set the message
  // The fp32 staging copies of the tensor-core SYRK are a speed-up, not a requirement: if they do not
  // fit the built-in array, plan this call without them (captures are then gathered in place).
  struct StageGuard {
    bool on = false;
    ~StageGuard() { if (on) set_stage_override(-1); }
  } guard;
  if (need > (size_t)KFB_STATIC_SCRATCH_BYTES) {
    set_stage_override(0);
    guard.on = true;
    need = factor_workspace_bytes(jobs, n_jobs, ctx);
    if (need == 0) return KFB_ERR_INVALID;
  }
  if (need > (size_t)KFB_STATIC_SCRATCH_BYTES)
    KFB_FAIL(KFB_ERR_WORKSPACE,
             "kfb_factor_update_static: these %d job(s) need %zu bytes of scratch but the built-in array holds "
             "%zu; split the call or use kfb_factor_update with a caller workspace",
             n_jobs, need, (size_t)KFB_STATIC_SCRATCH_BYTES);
  void* ws = nullptr;
  rc = static_scratch(device, &ws);   // first call on a device resolves the symbol address
  if (rc) return rc;
  // the shared factor-update body clears the accumulators it uses: whatever the previous call left
  // in the array is gone before anything is accumulated
  return factor_update_with(jobs, n_jobs, alpha, ctx, ws, (size_t)KFB_STATIC_SCRATCH_BYTES,
                            "kfb_factor_update_static", (cudaStream_t)stream_);
}

}  // extern "C"
