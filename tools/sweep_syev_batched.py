"""cusolverDnXsyevBatched time as a function of (n, batch count), next to the single-matrix syevd time,
to calibrate the cost model kfb_syevd_batch uses to split factors between the batched stream and the
syevd lanes."""
import ctypes as C
import time

import torch

lib = C.CDLL('libcusolver.so.11')
DEV = 'cuda:0'
CUDA_R_32F = 0
VEC, UPPER = 1, 1


def main():
    torch.zeros(1, device=DEV)
    h = C.c_void_p()
    assert lib.cusolverDnCreate(C.byref(h)) == 0
    params = C.c_void_p()
    assert lib.cusolverDnCreateParams(C.byref(params)) == 0
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.cusolverDnSetStream(h, C.c_void_p(stream)) == 0
    import sys
    sizes = [int(a) for a in sys.argv[1:]] or [64, 128, 256, 512, 576, 1024, 1152]
    for n in sizes:
        line = 'n=%5d ' % n
        # single syevd
        x = torch.randn(2 * n, n, device=DEV)
        ref1 = (x.t() @ x / (2 * n)).contiguous()
        A1 = ref1.clone()
        W1 = torch.empty(n, device=DEV)
        info = torch.zeros(64, dtype=torch.int32, device=DEV)
        lwork = C.c_int()
        assert lib.cusolverDnSsyevd_bufferSize(h, VEC, UPPER, n, C.c_void_p(A1.data_ptr()), n, C.c_void_p(W1.data_ptr()),
                                               C.byref(lwork)) == 0
        work = torch.empty(lwork.value, device=DEV)

        def single():
            return lib.cusolverDnSsyevd(h, VEC, UPPER, n, C.c_void_p(A1.data_ptr()), n, C.c_void_p(W1.data_ptr()),
                                        C.c_void_p(work.data_ptr()), lwork, C.c_void_p(info.data_ptr()))
        best = 1e9
        for _ in range(3):
            A1.copy_(ref1)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            single()
            torch.cuda.synchronize()
            best = min(best, (time.perf_counter() - t0) * 1e3)
        line += 'syevd %6.2f ms |' % best
        for batch in (1, 2, 4, 8, 16, 32):
            if (n >= 1024 and batch > 8) or (n >= 2048 and batch > 4) or (n >= 4608 and batch > 2):
                continue
            x = torch.randn(batch, 2 * n, n, device=DEV)
            ref = (x.transpose(1, 2) @ x / (2 * n)).contiguous()
            A = ref.clone()
            W = torch.empty(batch, n, device=DEV)
            dbytes, hbytes = C.c_size_t(), C.c_size_t()
            rc = lib.cusolverDnXsyevBatched_bufferSize(h, params, VEC, UPPER, C.c_int64(n), CUDA_R_32F,
                                                       C.c_void_p(A.data_ptr()), C.c_int64(n), CUDA_R_32F,
                                                       C.c_void_p(W.data_ptr()), CUDA_R_32F, C.byref(dbytes),
                                                       C.byref(hbytes), C.c_int64(batch))
            if rc != 0:
                line += ' b%d: rc%d' % (batch, rc)
                continue
            dbuf = torch.empty(max(dbytes.value, 1), dtype=torch.uint8, device=DEV)
            hbuf = (C.c_char * max(hbytes.value, 1))()
            db, hb = dbytes.value, hbytes.value

            def run():
                return lib.cusolverDnXsyevBatched(h, params, VEC, UPPER, C.c_int64(n), CUDA_R_32F,
                                                  C.c_void_p(A.data_ptr()), C.c_int64(n), CUDA_R_32F,
                                                  C.c_void_p(W.data_ptr()), CUDA_R_32F, C.c_void_p(dbuf.data_ptr()),
                                                  C.c_size_t(db), hbuf, C.c_size_t(hb), C.c_void_p(info.data_ptr()),
                                                  C.c_int64(batch))
            best = 1e9
            for _ in range(3):
                A.copy_(ref)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                run()
                torch.cuda.synchronize()
                best = min(best, (time.perf_counter() - t0) * 1e3)
            line += ' b%d %6.2f' % (batch, best)
        print(line, flush=True)


if __name__ == '__main__':
    main()
