"""Experiment: does launching each syevd lane through its own green context (cuGreenCtxCreate, a
slice of the SMs each) lift the aggregate throughput of concurrent cusolverDnSsyevd calls?  With
ordinary streams two host threads reach ~1.33x and more threads add nothing (one context's launch
path serialises)."""
import ctypes as C
import sys
import threading
import time

import torch
from cuda.bindings import driver as drv

lib = C.CDLL('libcusolver.so.11')
DEV = 'cuda:0'
VEC, UPPER = 1, 1


def chk(res):
    err = res[0]
    if int(err) != 0:
        raise RuntimeError('driver call failed: %s' % (err,))
    return res[1] if len(res) == 2 else res[1:]


class Lane(object):
    def __init__(self, stream_int, n, count):
        self.h = C.c_void_p()
        assert lib.cusolverDnCreate(C.byref(self.h)) == 0
        assert lib.cusolverDnSetStream(self.h, C.c_void_p(stream_int)) == 0
        self.n, self.count = n, count
        x = torch.randn(2 * n, n, device=DEV)
        self.ref = (x.t() @ x / (2 * n)).contiguous()
        self.A = [self.ref.clone() for _ in range(count)]
        self.W = torch.empty(n, device=DEV)
        self.info = torch.zeros(1, dtype=torch.int32, device=DEV)
        lwork = C.c_int()
        assert lib.cusolverDnSsyevd_bufferSize(self.h, VEC, UPPER, n, C.c_void_p(self.A[0].data_ptr()), n,
                                               C.c_void_p(self.W.data_ptr()), C.byref(lwork)) == 0
        self.lwork = lwork
        self.work = torch.empty(lwork.value, device=DEV)

    def run(self):
        for a in self.A:
            rc = lib.cusolverDnSsyevd(self.h, VEC, UPPER, self.n, C.c_void_p(a.data_ptr()), self.n,
                                      C.c_void_p(self.W.data_ptr()), C.c_void_p(self.work.data_ptr()), self.lwork,
                                      C.c_void_p(self.info.data_ptr()))
            assert rc == 0, rc


def timed(lanes):
    for l in lanes:
        for a in l.A:
            a.copy_(l.ref)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ths = [threading.Thread(target=l.run) for l in lanes]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    total = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    torch.zeros(1, device=DEV)
    torch.cuda.synchronize()
    # ordinary streams
    for k in (1, 2, 4):
        streams = [torch.cuda.Stream() for _ in range(k)]
        lanes = [Lane(s.cuda_stream, n, total // k) for s in streams]
        timed(lanes)
        print('n=%d  %d x syevd on %d ordinary stream(s)/thread(s): %8.1f ms' % (n, total, k, timed(lanes)), flush=True)
    # green contexts
    dev = chk(drv.cuDeviceGet(0))
    sm_res = chk(drv.cuDeviceGetDevResource(dev, drv.CUdevResourceType.CU_DEV_RESOURCE_TYPE_SM))
    print('device SM resource: %d SMs' % sm_res.sm.smCount, flush=True)
    for k, min_sm in ((4, 32), (4, 16), (6, 16), (8, 16), (8, 8)):
        groups, nb, rem = chk(drv.cuDevSmResourceSplitByCount(k, sm_res, 0, min_sm))
        print('split into %d groups of %s SMs (+%d left)' % (nb, [g.sm.smCount for g in groups[:nb]], rem.sm.smCount), flush=True)
        lanes, keep = [], []
        for g in groups[:nb]:
            desc = chk(drv.cuDevResourceGenerateDesc([g], 1))
            gctx = chk(drv.cuGreenCtxCreate(desc, dev, int(drv.CUgreenCtxCreate_flags.CU_GREEN_CTX_DEFAULT_STREAM)))
            st = chk(drv.cuGreenCtxStreamCreate(gctx, int(drv.CUstream_flags.CU_STREAM_NON_BLOCKING), 0))
            keep.append((gctx, st))
            lanes.append(Lane(int(st), n, total // nb))
        timed(lanes)
        print('n=%d  %d x syevd on %d green-context stream(s)/thread(s): %8.1f ms' % (n, total, nb, timed(lanes)), flush=True)
        if k == 4 and min_sm == 32:
            # event interop: an event of the primary context recorded on / waited by a green-context stream
            try:
                ev = torch.cuda.Event()
                gst = torch.cuda.ExternalStream(int(keep[0][1]))
                ev.record(torch.cuda.current_stream())
                gst.wait_event(ev)
                ev2 = torch.cuda.Event()
                ev2.record(gst)
                torch.cuda.current_stream().wait_event(ev2)
                torch.cuda.synchronize()
                print('event interop (primary-context events on green-context streams): ok', flush=True)
            except Exception as e:
                print('event interop failed: %r' % (e,), flush=True)


if __name__ == '__main__':
    main()
