"""Experiment: cusolverDnXsyevBatched (uniform-size batches) vs looped syevd, ResNet-50 factor sizes."""
import ctypes as C
import time

import torch

lib = C.CDLL('libcusolver.so.11')
DEV = 'cuda:0'
CUDA_R_32F = 0
VEC, UPPER = 1, 1   # CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER


def main():
    torch.zeros(1, device=DEV)
    h = C.c_void_p()
    assert lib.cusolverDnCreate(C.byref(h)) == 0
    params = C.c_void_p()
    assert lib.cusolverDnCreateParams(C.byref(params)) == 0
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.cusolverDnSetStream(h, C.c_void_p(stream)) == 0
    total = 0.0
    for n, batch in [(64, 22), (128, 12), (256, 30), (512, 20), (576, 3), (1024, 19), (1152, 4), (2048, 7),
                     (2304, 6), (4608, 3)]:
        x = torch.randn(batch, 2 * n, n, device=DEV)
        A = (x.transpose(1, 2) @ x / (2 * n)).contiguous()
        ref = A.clone()
        W = torch.empty(batch, n, device=DEV)
        info = torch.zeros(batch, dtype=torch.int32, device=DEV)
        dbytes, hbytes = C.c_size_t(), C.c_size_t()
        rc = lib.cusolverDnXsyevBatched_bufferSize(h, params, VEC, UPPER, C.c_int64(n), CUDA_R_32F,
                                                   C.c_void_p(A.data_ptr()), C.c_int64(n), CUDA_R_32F,
                                                   C.c_void_p(W.data_ptr()), CUDA_R_32F, C.byref(dbytes),
                                                   C.byref(hbytes), C.c_int64(batch))
        if rc != 0:
            print('n=%d bufferSize rc=%d' % (n, rc))
            continue
        dbuf = torch.empty(max(dbytes.value, 1), dtype=torch.uint8, device=DEV)
        hbuf = (C.c_char * max(hbytes.value, 1))()

        def run():
            return lib.cusolverDnXsyevBatched(h, params, VEC, UPPER, C.c_int64(n), CUDA_R_32F,
                                              C.c_void_p(A.data_ptr()), C.c_int64(n), CUDA_R_32F,
                                              C.c_void_p(W.data_ptr()), CUDA_R_32F, C.c_void_p(dbuf.data_ptr()),
                                              C.c_size_t(dbytes.value), hbuf, C.c_size_t(hbytes.value),
                                              C.c_void_p(info.data_ptr()), C.c_int64(batch))
        rc = run()
        torch.cuda.synchronize()
        ok = rc == 0 and int(info.abs().max()) == 0
        # check: column-major eigenvectors -> rows of the row-major view
        E = A[0].double()
        recon = E.t() @ torch.diag(W[0].double()) @ E
        err = (recon - ref[0].double()).abs().max().item() / ref[0].abs().max().item()
        A.copy_(ref)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        run()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        total += ms
        print('n=%5d batch=%3d  rc=%d ok=%s  recon err %.1e  %8.2f ms (%.2f ms per matrix) dev ws %.1f MB host ws %d B' %
              (n, batch, rc, ok, err, ms, ms / batch, dbytes.value / 1e6, hbytes.value))
    print('total %.1f ms' % total)


if __name__ == '__main__':
    main()
