"""Time cuSOLVER syevd through the C ABI per factor size, alone and as the ResNet-50 batch."""
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


def timed(fn, reps=2):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


def main():
    ops = N.get_ops()
    ops._enter(torch.zeros(1, device=DEV))
    for d in (() if '--batch-only' in sys.argv else (64, 128, 256, 512, 576, 1024, 1152, 2048, 2304, 4608)):
        job = make(d)
        ms = timed(lambda: ops.syevd(*job))
        ms_t = timed(lambda: torch.linalg.eigh(job[0]))
        print('d=%5d  kfb_syevd %8.2f ms   torch.linalg.eigh %8.2f ms' % (d, ms, ms_t))
    # ResNet-50 factor sizes (A and G of all 54 layers)
    sys.path.insert(0, os.path.join(ROOT, 'tests', 'golden'))
    import models
    dims = [x for pair in models.resnet_factor_dims('resnet50') for x in pair]
    for a in sys.argv:
        if a.startswith('--rank='):          # the factors the reference's allocator gives rank r of W: --rank=r/W
            from distributed_kfac_pytorch_b200 import utils
            r, w = (int(v) for v in a.split('=')[1].split('/'))
            pairs = models.resnet_factor_dims('resnet50')
            locs = utils.load_balance(w, [x ** 3 + y ** 3 for x, y in pairs])
            dims = [v for pair, l in zip(pairs, locs) if l == r for v in pair]
    jobs = [make(d) for d in dims]
    print('resnet50: %d factors, sum d^3 = %.3g' % (len(dims), sum(float(d) ** 3 for d in dims)))
    if '--batch-only' not in sys.argv:
        print('  one at a time: %.1f ms' % timed(lambda: [ops.syevd(*j) for j in jobs], reps=1))
    print('  syevd_batch  : %.1f ms' % timed(lambda: ops.syevd_batch(jobs), reps=2))


if __name__ == '__main__':
    main()
