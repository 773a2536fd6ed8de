"""Experiment (round 2, second): (1) syevd time per factor size and lane width (table for the lane
planner), (2) heterogeneous SM partitions built by SEQUENTIAL splits (split off w1 SMs, split the
remainder, ...) on the ResNet-50 mix, with a two-ended queue: the widest lane takes the largest job
left, the narrow lanes the smallest.
"""
import ctypes as C
import json
import sys
import threading
import time

import torch
from cuda.bindings import driver as drv

lib = C.CDLL('libcusolver.so.11')
DEV = 'cuda:0'
VEC, UPPER = 1, 1
MIX = {'n1': [4608] * 3 + [2304] * 6 + [2048] * 7 + [1152] * 4 + [576] * 3,
       'n1all': [4608] * 3 + [2304] * 6 + [2048] * 7 + [1152] * 4 + [576] * 3 + [1024] * 19 + [512] * 19,
       'n4': [4608, 2304, 2304, 2048, 2048, 1152, 1024, 1024, 1024, 1024, 1024, 512, 512, 512, 512, 512, 576]}


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
    def __init__(self, stream_int, width):
        self.h = C.c_void_p()
        assert lib.cusolverDnCreate(C.byref(self.h)) == 0
        assert lib.cusolverDnSetStream(self.h, C.c_void_p(stream_int)) == 0
        self.width = width
        self.work = {}
        self.W = torch.empty(8192, device=DEV)
        self.info = torch.zeros(1, dtype=torch.int32, device=DEV)

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


def run(lanes, sizes, label, reps=2):
    jobs = [(n, ref_matrix(n).clone()) for n in sorted(sizes, reverse=True)]
    for l in lanes:
        for n, a in jobs:
            l.prepare(n, a)
    widest = max(l.width for l in lanes)
    for rep in range(reps):
        for n, a in jobs:
            a.copy_(ref_matrix(n))
        queue = list(jobs)
        lock = threading.Lock()
        torch.cuda.synchronize()
        t0 = time.perf_counter()

        def body(lane):
            while True:
                with lock:
                    if not queue:
                        return
                    pick = queue.pop(0) if lane.width == widest else queue.pop()
                lane.solve(*pick)

        ths = [threading.Thread(target=body, args=(l,)) for l in lanes]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
    if label:
        print('%-64s %8.1f ms' % (label, ms), flush=True)
    return ms


KEEP = []


def green_lanes(dev, sm_res, widths):
    """Sequential splits: peel one group of >= w SMs off the remainder for every requested width."""
    lanes, got = [], []
    res = sm_res
    for w in widths:
        groups, nb, rem = chk(drv.cuDevSmResourceSplitByCount(1, res, 0, w))
        g = groups[0]
        desc = chk(drv.cuDevResourceGenerateDesc([g], 1))
        gctx = chk(drv.cuGreenCtxCreate(desc, dev, int(drv.CUgreenCtxCreate_flags.CU_GREEN_CTX_DEFAULT_STREAM)))
        st = chk(drv.cuGreenCtxStreamCreate(gctx, int(drv.CUstream_flags.CU_STREAM_NON_BLOCKING), 0))
        KEEP.append((gctx, st))
        lanes.append(Lane(int(st), g.sm.smCount))
        got.append(g.sm.smCount)
        res = rem
    return lanes, got, res.sm.smCount


def main():
    torch.zeros(1, device=DEV)
    torch.cuda.synchronize()
    dev = chk(drv.cuDeviceGet(0))
    sm_res = chk(drv.cuDeviceGetDevResource(dev, drv.CUdevResourceType.CU_DEV_RESOURCE_TYPE_SM))
    which = sys.argv[1] if len(sys.argv) > 1 else 'all'

    if which in ('all', 'table'):
        sizes = [64, 128, 256, 512, 576, 1024, 1152, 2048, 2304, 4608]
        table = {}
        s = torch.cuda.Stream()
        full = Lane(s.cuda_stream, 148)
        table['148'] = {str(n): round(run([full], [n], None), 2) for n in sizes}
        print('width 148 (ordinary stream):', table['148'], flush=True)
        for w in (120, 96, 72, 64, 48, 32, 24, 16, 8):
            t0 = time.perf_counter()
            lanes, got, left = green_lanes(dev, sm_res, [w])
            create_ms = (time.perf_counter() - t0) * 1e3
            table[str(got[0])] = {str(n): round(run(lanes, [n], None), 2) for n in sizes}
            print('width %d (asked %d, %d left, context created in %.0f ms):' % (got[0], w, left, create_ms),
                  table[str(got[0])], flush=True)
        print('TABLE ' + json.dumps(table), flush=True)

    if which in ('all', 'mix'):
        for mix in ('n1', 'n4'):
            sizes = MIX[mix]
            s = torch.cuda.Stream()
            run([Lane(s.cuda_stream, 148)], sizes, '%s: one ordinary stream' % mix)
            for widths in ([48, 48, 48], [32, 32, 32, 32], [96, 24, 24], [72, 24, 24, 24], [64, 40, 40],
                           [64, 32, 24, 24], [56, 32, 32, 24], [64, 24, 24, 16, 16], [48, 32, 32, 32],
                           [40, 32, 24, 24, 24], [24] * 6, [64, 16, 16, 16, 16, 16], [80, 32, 32],
                           [48, 24, 24, 24, 24]):
                try:
                    lanes, got, left = green_lanes(dev, sm_res, widths)
                    run(lanes, sizes, '%s: green lanes %s (+%d idle)' % (mix, got, left))
                except Exception as e:
                    print('%s: lanes %s failed: %r' % (mix, widths, e), flush=True)


if __name__ == '__main__':
    main()
