"""Is the precondition GEMM's operand delivery (45-53 bytes per clock and SM) a DRAM effect?  One stage-1-shaped
problem whose operands fit in L2 (A 4608 x 4608 = 85 MB, B 1024 x 4608 = 19 MB; 72 pair tiles = one wave), timed
with the operands in L2 (back-to-back repeats) and with L2 flushed before every call (a 256 MB write)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from distributed_kfac_pytorch_b200 import _native as N  # noqa: E402
from distributed_kfac_pytorch_b200.layers.base import alloc2d  # noqa: E402


def main():
    dev = 'cuda:0'
    ops = N.get_ops()
    ops._enter(torch.zeros(1, device=dev))
    M, Nn, K = 4608, 1024, 4608
    _, a = alloc2d(M, K, dev, zero=True); a.normal_()
    _, b = alloc2d(Nn, K, dev, zero=True); b.normal_()
    _, d = alloc2d(M, Nn, dev, zero=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(3):
        ops.gemm_tn(a, b, d)
    torch.cuda.synchronize()
    # the operands stay in L2: ten calls back to back in one event window (launch latency hidden after the first)
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            ops.gemm_tn(a, b, d)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / 10)
    ts.sort()
    print('L2-resident  median %.1f us = %.0f TFLOP/s' % (ts[2] * 1e3, 2.0 * M * Nn * K / ts[2] / 1e9))
    # L2 flushed by a 256 MB write queued right before each call (which also hides the launch latency)
    ts = []
    for _ in range(20):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.gemm_tn(a, b, d)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    print('L2 flushed   median %.1f us = %.0f TFLOP/s' % (ts[10] * 1e3, 2.0 * M * Nn * K / ts[10] / 1e9))


if __name__ == '__main__':
    main()
