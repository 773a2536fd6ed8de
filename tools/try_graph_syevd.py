"""Experiment: can cuSOLVER syevd (through kfb_syevd) be captured in a CUDA graph, and how fast do
replays run, alone and 8 at a time on separate streams?"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from distributed_kfac_pytorch_b200 import _native as N  # noqa: E402
from distributed_kfac_pytorch_b200.layers.base import alloc2d  # noqa: E402

DEV = 'cuda:0'


def make(d):
    x = torch.randn(2 * d, d, device=DEV)
    f = x.t() @ x / (2 * d)
    return (f.contiguous(), alloc2d(d, d, DEV)[1], alloc2d(d, d, DEV)[1], torch.empty(d, device=DEV),
            torch.zeros(1, dtype=torch.int32, device=DEV))


def main():
    ops = N.get_ops()
    ops._enter(torch.zeros(1, device=DEV))
    lib = ops.lib
    for d in (256, 1024, 2304):
        jobs = [make(d) for _ in range(8)]
        wss = [torch.empty(lib.kfb_syevd_workspace_bytes(d) + 1024, dtype=torch.uint8, device=DEV) for _ in range(8)]
        for j in jobs:
            ops.syevd(*j)
        torch.cuda.synchronize()
        ref_w = jobs[0][3].clone()
        graphs = []
        t0 = time.perf_counter()
        try:
            for j, ws in zip(jobs, wss):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    stream = torch.cuda.current_stream().cuda_stream
                    f, E, Q, w, info = j
                    rc = lib.kfb_syevd(f.data_ptr(), f.stride(0), d, E.data_ptr(), E.stride(0), Q.data_ptr(),
                                       Q.stride(0), w.data_ptr(), ws.data_ptr(), ws.numel(), info.data_ptr(), stream)
                    assert rc == 0, lib.kfb_last_error()
                graphs.append(g)
        except Exception as e:  # noqa
            print('d=%d capture FAILED: %s' % (d, str(e)[:300]))
            return
        torch.cuda.synchronize()
        print('d=%d captured 8 graphs in %.2f s' % (d, time.perf_counter() - t0))
        jobs[0][3].zero_()
        graphs[0].replay()
        torch.cuda.synchronize()
        print('   replay reproduces eigenvalues:', torch.allclose(jobs[0][3], ref_w, rtol=1e-4, atol=1e-6))
        t0 = time.perf_counter()
        for _ in range(3):
            graphs[0].replay()
        torch.cuda.synchronize()
        one = (time.perf_counter() - t0) / 3 * 1e3
        streams = [torch.cuda.Stream() for _ in range(8)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for g, s in zip(graphs, streams):
            with torch.cuda.stream(s):
                g.replay()
        torch.cuda.synchronize()
        eight = (time.perf_counter() - t0) * 1e3
        t0 = time.perf_counter()
        ops.syevd(*jobs[0])
        torch.cuda.synchronize()
        eager = (time.perf_counter() - t0) * 1e3
        print('   eager %.2f ms | graph replay %.2f ms | 8 graphs on 8 streams %.2f ms (%.2f ms each)' %
              (eager, one, eight, eight / 8))


if __name__ == '__main__':
    main()
