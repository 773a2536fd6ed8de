// Library context: error strings, per-device state, mode switches.
#include <cusolverDn.h>
#include <stdarg.h>

#include <atomic>
#include <mutex>

#include <cstdlib>

#include "kfb_common.cuh"

namespace kfb {

static thread_local char g_err[1024] = "";
static thread_local Context* g_ctx = nullptr;
static Context g_contexts[64];
static std::mutex g_mu;
// KFB_MMA_MODE=0/1/2 chooses the initial mode (kfb_set_mma_mode changes it later)
static int initial_mma_mode() {
  const char* e = getenv("KFB_MMA_MODE");
  const int m = e ? atoi(e) : 1;
  return m >= 0 && m <= 2 ? m : 1;
}
static int g_mma_mode = initial_mma_mode();
static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

Context* current_context() { return g_ctx; }

int require_context(Context** out) {
  if (!g_ctx) KFB_FAIL(KFB_ERR_INVALID, "kfb_init(device) has not been called on this thread");
  *out = g_ctx;
  return KFB_OK;
}

int mma_mode() { return g_mma_mode; }
int syrk_pair_mode() {
  const char* e = getenv("KFB_SYRK_2CTA");
  return e ? atoi(e) : 0;
}
static thread_local int g_stage_override = -1;
void set_stage_override(int v) { g_stage_override = v; }
int stage_mode() {
  if (g_stage_override >= 0) return g_stage_override;
  const char* e = getenv("KFB_SYRK_STAGE");
  return !(e && e[0] == '0');
}
int tma_mode() {
  const char* e = getenv("KFB_SYRK_TMA");
  return !(e && e[0] == '0');
}

static int init_device(int device) {
  KFB_REQUIRE(device >= 0 && device < 64, "device index %d out of range", device);
  Context& c = g_contexts[device];
  if (c.device == device) {
    g_ctx = &c;
    return KFB_OK;
  }
  std::lock_guard<std::mutex> lock(g_mu);
  if (c.device != device) {
    int count = 0;
    KFB_CUDA(cudaGetDeviceCount(&count));
    KFB_REQUIRE(device < count, "device %d does not exist (%d visible)", device, count);
    KFB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    KFB_CUDA(cudaGetDeviceProperties(&prop, device));
    c.num_sms = prop.multiProcessorCount;
    c.cc_major = prop.major;
    c.cc_minor = prop.minor;
    if (prop.major != 10)
      KFB_FAIL(KFB_ERR_UNSUPPORTED,
               "libkfac_b200 is built for sm_100a only; device %d is sm_%d%d", device, prop.major,
               prop.minor);
    cusolverDnHandle_t h = nullptr;
    cusolverStatus_t st = cusolverDnCreate(&h);
    if (st != CUSOLVER_STATUS_SUCCESS) KFB_FAIL(KFB_ERR_CUSOLVER, "cusolverDnCreate failed (%d)", (int)st);
    c.cusolver = h;
    // TMA descriptors are encoded by the driver; fetch the entry point through the runtime so
    // the library has no link-time dependency on libcuda.so
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) c.encode_tiled = fn;
    (void)cudaGetLastError();
    c.umma_ok = c.encode_tiled != nullptr;
    c.device = device;
  }
  g_ctx = &c;
  return KFB_OK;
}

}  // namespace kfb

extern "C" {

int kfb_version(void) { return KFB_ABI_VERSION; }

const char* kfb_last_error(void) { return kfb::get_error(); }

int kfb_init(int device) { return kfb::init_device(device); }

int kfb_set_mma_mode(int mode) {
  KFB_REQUIRE(mode >= 0 && mode <= 2, "mma mode must be 0 (SIMT), 1 (tcgen05 TF32) or 2 (tcgen05, 3xTF32 precondition chain), got %d", mode);
  kfb::g_mma_mode = mode;
  return KFB_OK;
}

int kfb_get_mma_mode(void) { return kfb::g_mma_mode; }

unsigned long long kfb_launch_count(void) { return kfb::g_launches.load(); }

}  // extern "C"
