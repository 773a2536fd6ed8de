"""Experiment: the 64-bit generic cusolverDnXsyevd vs the legacy cusolverDnSsyevd (what kfb_syevd
calls) per factor size."""
import ctypes as C
import time

import torch

lib = C.CDLL('libcusolver.so.11')
DEV = 'cuda:0'
CUDA_R_32F = 0
VEC, UPPER, RANGE_ALL = 1, 1, 1001   # CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, CUSOLVER_EIG_RANGE_ALL


def timed(fn, restore, reps=2):
    restore()
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        restore()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        best = min(best, (time.perf_counter() - t0) * 1e3)
    return best


def main():
    torch.zeros(1, device=DEV)
    h = C.c_void_p()
    assert lib.cusolverDnCreate(C.byref(h)) == 0
    params = C.c_void_p()
    assert lib.cusolverDnCreateParams(C.byref(params)) == 0
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.cusolverDnSetStream(h, C.c_void_p(stream)) == 0
    for n in (512, 1024, 2048, 2304, 4608):
        x = torch.randn(2 * n, n, device=DEV)
        ref = (x.t() @ x / (2 * n)).contiguous()
        A = ref.clone()
        W = torch.empty(n, device=DEV)
        info = torch.zeros(1, dtype=torch.int32, device=DEV)

        def restore():
            A.copy_(ref)

        # legacy
        lwork = C.c_int()
        assert lib.cusolverDnSsyevd_bufferSize(h, VEC, UPPER, n, C.c_void_p(A.data_ptr()), n, C.c_void_p(W.data_ptr()),
                                               C.byref(lwork)) == 0
        work = torch.empty(lwork.value, device=DEV)
        t_legacy = timed(lambda: lib.cusolverDnSsyevd(h, VEC, UPPER, n, C.c_void_p(A.data_ptr()), n,
                                                      C.c_void_p(W.data_ptr()), C.c_void_p(work.data_ptr()), lwork,
                                                      C.c_void_p(info.data_ptr())), restore)
        # generic 64-bit
        dbytes, hbytes = C.c_size_t(), C.c_size_t()
        rc = lib.cusolverDnXsyevd_bufferSize(h, params, VEC, UPPER, C.c_int64(n), CUDA_R_32F, C.c_void_p(A.data_ptr()),
                                             C.c_int64(n), CUDA_R_32F, C.c_void_p(W.data_ptr()), CUDA_R_32F,
                                             C.byref(dbytes), C.byref(hbytes))
        t_x = float('nan')
        err = float('nan')
        if rc == 0:
            dbuf = torch.empty(max(dbytes.value, 1), dtype=torch.uint8, device=DEV)
            hbuf = (C.c_char * max(hbytes.value, 1))()
            t_x = timed(lambda: lib.cusolverDnXsyevd(h, params, VEC, UPPER, C.c_int64(n), CUDA_R_32F,
                                                     C.c_void_p(A.data_ptr()), C.c_int64(n), CUDA_R_32F,
                                                     C.c_void_p(W.data_ptr()), CUDA_R_32F, C.c_void_p(dbuf.data_ptr()),
                                                     C.c_size_t(dbytes.value), hbuf, C.c_size_t(hbytes.value),
                                                     C.c_void_p(info.data_ptr())), restore)
            E = A.double()
            recon = E.t() @ torch.diag(W.double()) @ E
            err = (recon - ref.double()).abs().max().item() / ref.abs().max().item()
        t_dx = float('nan')
        print('n=%5d  Ssyevd %8.2f ms   Xsyevd %8.2f ms (recon err %.1e)   Xsyevdx(all) %8.2f ms' % (n, t_legacy, t_x, err, t_dx), flush=True)


if __name__ == '__main__':
    main()
