"""Experiment (round 2): lanes of DIFFERENT widths and OVERLAPPING green contexts for concurrent
cusolverDnSsyevd calls.  Workload = the single (non-batched) solves of ResNet-50 on one GPU:
3 x 4608, 6 x 2304, 7 x 2048, 4 x 1152, 3 x 576; lanes pull the next largest job from one queue.

  A  k disjoint equal green contexts (what round 1 shipped: 3 x 48 SMs)
  B  k green contexts that ALL span the same SMs (does the driver allow it? do launches still
     run in parallel?)
  C  one wide lane + narrow lanes (the 4608s on the wide one)
"""
import ctypes as C
import sys
import threading
import time

import torch
from cuda.bindings import driver as drv

lib = C.CDLL('libcusolver.so.11')
DEV = 'cuda:0'
VEC, UPPER = 1, 1
SIZES = [4608] * 3 + [2304] * 6 + [2048] * 7 + [1152] * 4 + [576] * 3
if len(sys.argv) > 1 and sys.argv[1] == 'n8':       # rank 0 of 8: one 4608 + one 512 and little else
    SIZES = [4608, 512, 256, 64]


def chk(res):
    err = res[0]
    if int(err) != 0:
        raise RuntimeError('driver call failed: %s' % (err,))
    return res[1] if len(res) == 2 else res[1:]


REF = {}


def ref_matrix(n):
    if n not in REF:
        x = torch.randn(2 * n, n, device=DEV)
        REF[n] = (x.t() @ x / (2 * n)).contiguous()
    return REF[n]


class Lane(object):
    def __init__(self, stream_int, accept=None):
        self.h = C.c_void_p()
        assert lib.cusolverDnCreate(C.byref(self.h)) == 0
        assert lib.cusolverDnSetStream(self.h, C.c_void_p(stream_int)) == 0
        self.accept = accept
        self.work = {}
        self.W = torch.empty(8192, device=DEV)
        self.info = torch.zeros(1, dtype=torch.int32, device=DEV)
        self.log = []

    def prepare(self, n, a):
        if n not in self.work:
            lwork = C.c_int()
            assert lib.cusolverDnSsyevd_bufferSize(self.h, VEC, UPPER, n, C.c_void_p(a.data_ptr()), n,
                                                   C.c_void_p(self.W.data_ptr()), C.byref(lwork)) == 0
            self.work[n] = (lwork, torch.empty(lwork.value, device=DEV))

    def solve(self, n, a):
        lwork, work = self.work[n]
        rc = lib.cusolverDnSsyevd(self.h, VEC, UPPER, n, C.c_void_p(a.data_ptr()), n, C.c_void_p(self.W.data_ptr()),
                                  C.c_void_p(work.data_ptr()), lwork, C.c_void_p(self.info.data_ptr()))
        assert rc == 0, rc


def run(lanes, label):
    jobs = [(n, ref_matrix(n).clone()) for n in sorted(SIZES, reverse=True)]
    for l in lanes:
        for n, a in jobs:
            l.prepare(n, a)
    for rep in range(2):
        for n, a in jobs:
            a.copy_(ref_matrix(n))
        queue = list(jobs)
        lock = threading.Lock()
        torch.cuda.synchronize()
        t0 = time.perf_counter()

        def body(lane):
            while True:
                with lock:
                    pick = None
                    for i, (n, a) in enumerate(queue):
                        if lane.accept is None or lane.accept(n):
                            pick = queue.pop(i)
                            break
                if pick is None:
                    return
                lane.solve(*pick)

        ths = [threading.Thread(target=body, args=(l,)) for l in lanes]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
    print('%-70s %8.1f ms' % (label, ms), flush=True)
    return ms


def green_stream(dev, resources):
    desc = chk(drv.cuDevResourceGenerateDesc(resources, len(resources)))
    gctx = chk(drv.cuGreenCtxCreate(desc, dev, int(drv.CUgreenCtxCreate_flags.CU_GREEN_CTX_DEFAULT_STREAM)))
    st = chk(drv.cuGreenCtxStreamCreate(gctx, int(drv.CUstream_flags.CU_STREAM_NON_BLOCKING), 0))
    return gctx, st


def main():
    torch.zeros(1, device=DEV)
    torch.cuda.synchronize()
    keep = []
    s = torch.cuda.Stream()
    run([Lane(s.cuda_stream)], 'one ordinary stream')
    dev = chk(drv.cuDeviceGet(0))
    sm_res = chk(drv.cuDeviceGetDevResource(dev, drv.CUdevResourceType.CU_DEV_RESOURCE_TYPE_SM))
    print('device SM resource: %d SMs' % sm_res.sm.smCount, flush=True)

    # A: disjoint equal lanes
    for k, per in ((3, 48), (4, 32)):
        try:
            groups, nb, rem = chk(drv.cuDevSmResourceSplitByCount(k, sm_res, 0, per))
            lanes = []
            for g in groups[:nb]:
                gctx, st = green_stream(dev, [g])
                keep.append((gctx, st))
                lanes.append(Lane(int(st)))
            run(lanes, 'A: %d disjoint green contexts x %d SMs' % (nb, groups[0].sm.smCount))
        except Exception as e:
            print('A k=%d failed: %r' % (k, e), flush=True)

    # B: overlapping - every context spans the whole device
    for k in (2, 3, 4, 6, 8):
        try:
            lanes = []
            for i in range(k):
                gctx, st = green_stream(dev, [sm_res])
                keep.append((gctx, st))
                lanes.append(Lane(int(st)))
            run(lanes, 'B: %d green contexts, each over all %d SMs' % (k, sm_res.sm.smCount))
        except Exception as e:
            print('B k=%d failed: %r' % (k, e), flush=True)

    # B2: overlapping, each over the same 128-SM group
    try:
        groups, nb, rem = chk(drv.cuDevSmResourceSplitByCount(1, sm_res, 0, 128))
        for k in (4, 8):
            lanes = []
            for i in range(k):
                gctx, st = green_stream(dev, [groups[0]])
                keep.append((gctx, st))
                lanes.append(Lane(int(st)))
            run(lanes, 'B2: %d green contexts, each over the same %d SMs' % (k, groups[0].sm.smCount))
    except Exception as e:
        print('B2 failed: %r' % (e,), flush=True)

    # C: one wide lane (built from several 8-SM groups) + narrow lanes
    try:
        groups, nb, rem = chk(drv.cuDevSmResourceSplitByCount(18, sm_res, 0, 8))
        print('split into %d groups of %d SMs (+%d left)' % (nb, groups[0].sm.smCount, rem.sm.smCount), flush=True)
        for wide, narrow in ((12, [3, 3]), (10, [4, 4]), (8, [5, 5]), (9, [3, 3, 3])):
            lanes = []
            gctx, st = green_stream(dev, list(groups[:wide]))
            keep.append((gctx, st))
            lanes.append(Lane(int(st), accept=lambda n: n >= 2304))
            off = wide
            for w in narrow:
                gctx, st = green_stream(dev, list(groups[off:off + w]))
                keep.append((gctx, st))
                lanes.append(Lane(int(st), accept=lambda n: n < 4608))
                off += w
            run(lanes, 'C: wide %d SMs (d>=2304) + narrow %s SMs (d<4608)' % (wide * 8, [w * 8 for w in narrow]))
    except Exception as e:
        print('C failed: %r' % (e,), flush=True)


if __name__ == '__main__':
    main()
