/*
 * kfac_b200.h — C ABI of libkfac_b200.so, the B200 (sm_100a) implementation of the
 * K-FAC preconditioner hot path of MLHPC/Distributed_KFAC_Pytorch.
 *
 * The reference is pure Python (no FFI of its own), so the "binding" a maintainer
 * adds is the ctypes stub shown in INTEGRATION.md; each entry point below cites the
 * reference function(s) it replaces as /root/reference file:line.
 *
 * Conventions
 *   - every function returns 0 on success, a kfb_status code otherwise;
 *     kfb_last_error() returns a thread-local message for the last failure;
 *   - all matrices are fp32, row-major, with an explicit row pitch `ld*` in ELEMENTS;
 *   - device pointers are plain `void*` / `float*` (tensor.data_ptr()); `stream` is a
 *     cudaStream_t passed as void* (0 = legacy default stream);
 *   - job arrays live in HOST memory and are consumed before the call returns;
 *   - nothing here allocates device memory: scratch comes from the caller, sized by
 *     the matching *_workspace_bytes() function;
 *   - calls are asynchronous with respect to the host unless stated otherwise.
 */
#ifndef KFAC_B200_H_
#define KFAC_B200_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define KFB_API __attribute__((visibility("default")))
#else
#define KFB_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  KFB_OK = 0,
  KFB_ERR_INVALID = 1,     /* bad argument (null pointer, negative size, pitch too small ...) */
  KFB_ERR_CUDA = 2,        /* a CUDA runtime/driver call failed; message has the cudaError string */
  KFB_ERR_UNSUPPORTED = 3, /* shape outside what the kernels support */
  KFB_ERR_CUSOLVER = 4,
  KFB_ERR_WORKSPACE = 5    /* caller-provided workspace too small */
} kfb_status;

#define KFB_ABI_VERSION 2

/* ---- library ------------------------------------------------------------------ */
KFB_API int kfb_version(void);
KFB_API const char* kfb_last_error(void);
/* Create the per-device context (cuSOLVER handle, SM count, TMA driver entry point) for
 * `device` and make it the context used by later calls from this thread. Idempotent. */
KFB_API int kfb_init(int device);
/* Which code path the GEMM-shaped kernels use: 0 = SIMT fp32 (exact fp32 FMA, the
 * first-correct path), 1 = tcgen05 TF32 tensor cores (default when available), 2 = as 1, but the
 * precondition chain (kfac/layers/base.py:459-470) forms every product as the error-compensated
 * "3xTF32" sum A_lo B_hi + A_hi B_lo + A_hi B_hi with fp32 accumulation: fp32-grade gradients at a
 * third of the tensor-core rate. */
KFB_API int kfb_set_mma_mode(int mode);
KFB_API int kfb_get_mma_mode(void);
/* Number of kernels of this library launched so far in this process (all devices); cuSOLVER's
 * internal kernels and memcpy/memset are not counted. Used by bench.py for "gpu_launches". */
KFB_API unsigned long long kfb_launch_count(void);

/* ---- factor computation + running average ------------------------------------------
 * Replaces Conv2dLayer._get_A_factor/_get_G_factor/_extract_patches (kfac/layers/conv.py:24-70),
 * LinearLayer / LinearMultiLayer._get_{A,G}_factor (kfac/layers/linear.py:12-59),
 * append_bias_ones/get_cov/reshape_data (kfac/layers/utils.py:4-43,107-124),
 * update_running_avg (kfac/layers/utils.py:164-178) and the identity seed of
 * KFACLayer.update_{A,G}_factor (kfac/layers/base.py:381-417).
 *
 * A factor is   F = sum_segments weight_s * X_s^T X_s   (upper triangle computed, mirrored),
 * followed by   state <- (state * alpha/(1-alpha) + F) * (1-alpha)    (skipped when alpha == 1).
 * X_s is never materialised; it is read through one of three layouts: */
enum {
  KFB_SEG_ROWS = 0,   /* X[r][f] = data[r*ld + f], r < n, f < c  (Linear a / g, any leading dims collapsed) */
  KFB_SEG_NCHW = 1,   /* X[(i,s)][f] = data[(i*c + f)*h*w + s], i < n, s < h*w  (conv grad_output; 1x1/s1 conv input) */
  KFB_SEG_IM2COL = 2  /* X[(i,oy,ox)][(f,ky,kx)] = data[i][f][oy*sh+ky-ph][ox*sw+kx-pw], 0 outside  (conv input) */
};

typedef struct kfb_segment {
  int32_t kind;
  int32_t has_bias;      /* 1: X has one extra trailing feature that is constant 1 (ROWS, IM2COL) */
  const void* data;
  int64_t n;             /* ROWS: row count; NCHW / IM2COL: batch size */
  int64_t c;             /* feature / channel count (without the bias feature) */
  int64_t h, w;          /* NCHW: spatial extent; IM2COL: INPUT height and width */
  int32_t kh, kw, sh, sw, ph, pw;   /* IM2COL kernel, stride, padding */
  int64_t ld;            /* ROWS: row pitch in elements */
  float weight;          /* e.g. 1/(N*S^3) for conv, 1/M for linear */
  int32_t dtype;         /* element type of `data`: KFB_F32 (0), KFB_F16 (1) or KFB_BF16 (2).  16-bit captures
                            (AMP autocast activations / output gradients) are multiplied as they are, by
                            16-bit tensor-core instructions with exact products (or widened to fp32 where
                            that path does not apply); accumulation and the factor stay fp32 */
  const float* dev_scale; /* NULL, or a DEVICE float that multiplies `weight` when the kernels run; 0 drops the
                            segment altogether (its data may then hold inf / nan).  Written by
                            kfb_capture_filter: the GradScaler handling of kfac/layers/base.py:397-407
                            without a host round trip */
} kfb_segment;
enum { KFB_F32 = 0, KFB_F16 = 1, KFB_BF16 = 2 };

typedef struct kfb_factor_job {
  int32_t d;                     /* factor dimension, = c*kh*kw + has_bias */
  int32_t n_segments;
  const kfb_segment* segments;   /* host array */
  float* state;                  /* d x d running average, pitch ld_state */
  int64_t ld_state;
  int32_t seed_identity;         /* 1: treat the old state as the identity (first update) */
  int32_t write_state;           /* 1: write the full (mirrored) square into state */
  float* tri_out;                /* NULL, or d(d+1)/2 floats: packed upper triangle (row by row) of the
                                    new state - a slot of the caller's allreduce fusion buffer */
  const float* dev_rescale;      /* NULL, or a DEVICE float r: r >= 0 multiplies F (rows captured / rows kept
                                    after kfb_capture_filter dropped segments); r < 0 means every segment was
                                    dropped: the running average is left as it is (base.py:409-410) */
} kfb_factor_job;

/* AMP captures: which output-gradient captures are usable, decided ON THE DEVICE.  Replaces the
 * unscale + `torch.isinf(g).any() or torch.isnan(g).any()` of KFACLayer.update_G_factor
 * (kfac/layers/base.py:397-407), which costs the reference one host synchronisation per capture.
 * For every capture c:  dev_scale[c] = all_finite(data_c) ? inv_scale_sq : 0, where inv_scale_sq is
 * 1 / s^2 with the loss scale s read from the device float `loss_scale` (NULL: host_inv_scale_sq is
 * used) - G is quadratic in g, so (g/s)^T (g/s) = g^T g / s^2.  Captures of one factor share a
 * `group`; for every group g: dev_rescale[g] = -1 if no capture survived, else rows(all) / rows(kept)
 * when `renormalise` (the reference concatenates the surviving captures and divides by THEIR row
 * count, kfac/layers/utils.py:107-124,36-41) or 1 (LinearMultiLayer: every call has its own divisor). */
typedef struct kfb_filter_job {
  const void* data; int64_t numel; int32_t dtype;   /* the capture */
  int32_t group;                                    /* index into the group arrays below */
  float rows;                                       /* rows of X this capture contributes */
  float host_inv_scale_sq;                          /* used when loss_scale == NULL */
  const float* loss_scale;                          /* device float (GradScaler._scale) or NULL */
  float* dev_scale;                                 /* out: device float, see kfb_segment.dev_scale */
} kfb_filter_job;
typedef struct kfb_filter_group {
  float* dev_rescale;                               /* out: device float, see kfb_factor_job.dev_rescale */
  int32_t renormalise;
  int32_t reserved;
} kfb_filter_group;
KFB_API int kfb_capture_filter(const kfb_filter_job* jobs, int32_t n_jobs, const kfb_filter_group* groups,
                               int32_t n_groups, void* stream);

/* ---- 16-bit storage of the inverses (inv_dtype = float16 / bfloat16, kfac/layers/base.py:435-441) ----
 * The eigendecomposition runs in fp32; its results are then ROUNDED to the 16-bit type, travel in 16
 * bits (halving the inverse broadcast) and are widened again for the fp32 / TF32 GEMM chain.
 * dst16[i] = narrow(src[i]) and dst[i] = widen(src16[i]) over flat arrays; `dtype` is KFB_F16 / KFB_BF16. */
KFB_API int kfb_narrow(const float* src, void* dst16, int64_t n, int32_t dtype, void* stream);
KFB_API int kfb_widen(const void* src16, float* dst, int64_t n, int32_t dtype, void* stream);

/* Both factor-update entry points plan on the host and, on every call, upload their descriptor tables
 * with cudaMemcpyAsync from ordinary (pageable) host memory owned by the library's plan cache and
 * clear their accumulators with cudaMemsetAsync before the kernels run. */
KFB_API size_t kfb_factor_workspace_bytes(const kfb_factor_job* jobs, int32_t n_jobs);
KFB_API int kfb_factor_update(const kfb_factor_job* jobs, int32_t n_jobs, float alpha,
                      void* workspace, size_t workspace_bytes, void* stream);

/* Split-K jobs accumulate into the workspace and are finished (scale, running average, mirror:
 * kfac/layers/utils.py:164-178) inside the SYRK by whichever work item of a 256 x 256 block arrives last;
 * that item also re-zeroes the block's accumulator and its arrival counter.  A caller that uses a
 * workspace for kfb_factor_update ONLY can say so: the library then relies on what it left there and
 * skips the per-call clear of those accumulators when the same set of jobs (same shapes) comes again on
 * a stream ordered after the previous call.  Anything else writing into a dedicated workspace between
 * two updates corrupts the second.  `release` withdraws the promise (call it before freeing the
 * memory).  Neither call touches the device. */
KFB_API int kfb_factor_workspace_dedicate(void* workspace, size_t workspace_bytes);
KFB_API int kfb_factor_workspace_release(void* workspace);

/* kfb_factor_update for callers that cannot allocate device memory.  No counterpart in the
 * reference (which allocates freely through torch); it exists for hosts that own all GPU memory
 * themselves or that may not call cudaMalloc once running.  Same jobs, same results, but the
 * scratch (descriptors + split-K accumulators) is a fixed KFB_STATIC_SCRATCH_BYTES array inside the
 * library: every GPU that loads the library holds that much memory until the process exits.
 *   - nothing is allocated or freed on the device, here or in any library it calls (planning uses
 *     ordinary host heap memory);
 *   - a call whose jobs need more than kfb_static_scratch_bytes() fails with KFB_ERR_WORKSPACE and
 *     a message giving both sizes - split the jobs over several calls, or use kfb_factor_update;
 *   - the first call on each device looks up the array's device address (with that device
 *     current); make that first call before starting any stream capture;
 *   - all calls on a device share the one array and it keeps the previous call's contents; each
 *     call clears the part it accumulates into, so back-to-back calls on ONE stream are fine, but
 *     two calls whose kernels can overlap on the GPU (different streams, different threads) race. */
#define KFB_STATIC_SCRATCH_BYTES (16u << 20)
KFB_API size_t kfb_static_scratch_bytes(void);
KFB_API int kfb_factor_update_static(const kfb_factor_job* jobs, int32_t n_jobs, float alpha, void* stream);

/* ---- packed-triangle fusion buffer ----------------------------------------------------
 * Replaces get_triu / fill_triu (kfac/layers/utils.py:126-162) and the deferred
 * `tensor /= world` of TorchBackend.sync (kfac/comm.py:254-262). */
typedef struct kfb_tri_job {
  float* tri;          /* d(d+1)/2 */
  float* state;        /* d x d, pitch ld_state */
  int64_t ld_state;
  int32_t d;
  int32_t reserved;
} kfb_tri_job;
/* state <- scale * unpack(tri), both triangles */
KFB_API int kfb_tri_unpack(const kfb_tri_job* jobs, int32_t n_jobs, float scale, void* stream);
/* tri <- upper triangle of state */
KFB_API int kfb_tri_pack(const kfb_tri_job* jobs, int32_t n_jobs, void* stream);

/* ---- eigendecomposition (the one library call: cuSOLVER syevd) ----------------------------
 * Replaces get_eigendecomp (kfac/layers/utils.py:45-74) as used by
 * KFACLayer._compute_factor_inverse (kfac/layers/base.py:432-441): eigenvalues are clipped
 * at 0.  Outputs: E (rows are eigenvectors, E = Q^T), Q (columns are eigenvectors, the
 * reference's state['QA'|'QG']) and w.  `factor` is not modified. `dev_info` is a device
 * int32 written by cuSOLVER (0 = converged); the call does not synchronise. */
KFB_API size_t kfb_syevd_workspace_bytes(int32_t d);
KFB_API int kfb_syevd(const float* factor, int64_t ld_factor, int32_t d,
              float* E, int64_t ld_e, float* Q, int64_t ld_q, float* w,
              void* workspace, size_t workspace_bytes, int32_t* dev_info, void* stream);
/* Many eigendecompositions at once, still all cuSOLVER: factors of the same size (small and medium
 * ones, where it measures faster) are solved together by cusolverDnXsyevBatched; the rest are
 * spread over internal worker lanes - one host thread, CUDA stream and cuSOLVER handle each.  Where
 * the driver has green contexts (CUDA >= 12.4) the lanes run on DISJOINT SM partitions of DIFFERENT
 * widths, planned per call from the job list with a measured (factor size x SMs) time table: large
 * factors need many SMs, small ones lose little on 8 or 16 (kfb_syevd_lane_plan shows the plan).  A
 * job list that one full-width stream serves best (e.g. a single large factor) gets exactly that.
 * Partitions are created on first use (a few ms per lane) and kept for the life of the process.
 * Under a profiler, with KFB_SYEVD_GREEN=0, or when the driver refuses a partition, the lanes are
 * two ordinary streams.  Everything starts after the work already queued on `stream` and `stream`
 * waits for all of it, so to the caller it behaves like one asynchronous call on `stream`.  One call
 * at a time per device (a second caller blocks until the first has queued its work).  Results match
 * kfb_syevd job by job (eigenvectors up to sign). */
typedef struct kfb_syevd_job {
  const float* factor; int64_t ld_factor;
  float* E; int64_t ld_e;
  float* Q; int64_t ld_q;
  float* w;
  int32_t* dev_info;
  int32_t d;
  int32_t reserved;
} kfb_syevd_job;
KFB_API size_t kfb_syevd_batch_workspace_bytes(const kfb_syevd_job* jobs, int32_t n_jobs);
/* The lane planner of kfb_syevd_batch on its own (host only, no GPU needed): given the sizes of the
 * factors that go to syevd, how the SMs of a `num_sms`-SM B200 would be partitioned into at most
 * `max_lanes` (<= 8) lanes and which lane each factor would run on, from the measured (size x SMs)
 * time table built into the library.  lane_of_job[n], lane_sms[8]; *modelled_ms = planned makespan.
 * The placement of the eigendecompositions on RANKS is the reference's allocator
 * (kfac/utils.py:169-196, bit-exact in utils.py); this is the placement inside one rank's GPU, which
 * the reference leaves to a serial loop (kfac/preconditioner.py:555-564). */
KFB_API int kfb_syevd_lane_plan(const int32_t* dims, int32_t n, int32_t num_sms, int32_t max_lanes,
                                int32_t* lane_of_job, int32_t* lane_sms, int32_t* n_lanes, double* modelled_ms);
KFB_API int kfb_syevd_batch(const kfb_syevd_job* jobs, int32_t n_jobs, void* workspace, size_t workspace_bytes,
                            void* stream);
/* Explicit damped inverse, out = (factor + damping I)^-1, by Cholesky factorisation (cuSOLVER potrf +
 * potri) - replaces get_inverse (kfac/layers/utils.py:76-96) for use_eigen_decomp=False.  `out` is
 * written as a full symmetric matrix; `factor` is not modified; `dev_info` as for kfb_syevd. */
KFB_API size_t kfb_cholesky_inverse_workspace_bytes(int32_t d);
KFB_API int kfb_cholesky_inverse(const float* factor, int64_t ld_factor, int32_t d, float damping,
                                 float* out, int64_t ld_out, void* workspace, size_t workspace_bytes,
                                 int32_t* dev_info, void* stream);
/* out <- in^T  (receivers of a broadcast Q rebuild E) */
KFB_API int kfb_transpose(const float* in, int64_t ld_in, int32_t rows, int32_t cols,
                  float* out, int64_t ld_out, void* stream);
/* out[i][j] = 1 / (dG[i]*dA[j] + damping)    (kfac/layers/base.py:305-306) */
KFB_API int kfb_outer_recip(const float* dG, const float* dA, int32_t dg, int32_t da, float damping,
                    float* out, int64_t ld_out, void* stream);

/* ---- preconditioning -------------------------------------------------------------------
 * Replaces KFACLayer.get_gradient (kfac/layers/base.py:310-319, conv.py:17-22),
 * _get_precondition_gradient_eigen (base.py:459-470) and the per-layer `.sum().item()` of
 * KFAC._compute_grad_scale (kfac/preconditioner.py:661-682).
 *
 *   gmat = [weight_grad.view(dg,-1) | bias_grad]
 *   out  = QG ((QG^T gmat QA) * dGdA) QA^T
 *   *vg_acc += sum(out * gmat)          (device scalar, NOT multiplied by lr^2)
 *
 * run as four grouped "TN" GEMM stages over all jobs (D[m][n] = sum_k A[m][k] B[n][k]):
 *   T = EA gmat^T ; V = (EG T^T) * dGdA ; T = QA V^T ; out = QG T^T.            */
typedef struct kfb_precond_job {
  int32_t da, dg;          /* da includes the bias column */
  int32_t has_bias;
  int32_t use_inverse;     /* 1: explicit-inverse mode, out = G_inv gmat A_inv (kfac/layers/base.py:472-475):
                              EA must point to A_inv and QG to G_inv (both symmetric); QA, EG, dGdA, V unused */
  const float* weight_grad;   /* contiguous [dg][da - has_bias] */
  const float* bias_grad;     /* [dg] or NULL */
  const float* EA; const float* QA; int64_t ld_a;   /* da x da */
  const float* EG; const float* QG; int64_t ld_g;   /* dg x dg */
  const float* dGdA; int64_t ld_dgda;               /* dg x da */
  float* gmat; float* V; float* out; int64_t ld_v;  /* dg x da each, same pitch */
  float* T; int64_t ld_t;                           /* da x dg */
} kfb_precond_job;
KFB_API int kfb_precondition(const kfb_precond_job* jobs, int32_t n_jobs, float* vg_acc, void* stream);

/* Plain grouped-GEMM utility on the same kernels: D[m][n] = sum_k A[m][k] * B[n][k] (D = A B^T),
 * optionally multiplied elementwise by Mul[m][n].  This is the building block of the chain above
 * and of the explicit-inverse mode (G_inv grad A_inv, kfac/layers/base.py:472-475). */
KFB_API int kfb_gemm_tn(const float* A, int64_t lda, const float* B, int64_t ldb, float* D, int64_t ldd,
                        int32_t M, int32_t N, int32_t K, const float* Mul, int64_t ldmul, void* stream);

/* *vg_acc += sum_jobs sum(v * [weight_grad | bias_grad])   — for layers whose preconditioned
 * gradient arrived by broadcast (MEM_OPT / HYBRID_OPT). */
typedef struct kfb_grad_job {
  float* weight_grad;      /* contiguous [dg][da - has_bias] */
  float* bias_grad;        /* [dg] or NULL */
  const float* v;          /* dg x da preconditioned gradient, pitch ld_v */
  int64_t ld_v;
  int32_t da, dg, has_bias;
  int32_t reserved;
} kfb_grad_job;
KFB_API int kfb_dot_accumulate(const kfb_grad_job* jobs, int32_t n_jobs, float* vg_acc, void* stream);

/* Replaces KFAC.update_gradients / KFACLayer.update_gradient / _set_*_grad
 * (kfac/preconditioner.py:587-590, kfac/layers/base.py:419-430,477-483) and the scalar part of
 * _compute_grad_scale: writes nu * v into weight_grad / bias_grad IN PLACE, where
 *   nu = min(1, sqrt(kl_clip / |*vg_acc * lr^2|))   if vg_acc != NULL and *vg_acc != 0,
 *   nu = fixed_scale                                otherwise.
 * No host synchronisation. */
KFB_API int kfb_scale_writeback(const kfb_grad_job* jobs, int32_t n_jobs, const float* vg_acc,
                        float lr, float kl_clip, float fixed_scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KFAC_B200_H_ */
