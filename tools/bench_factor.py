"""Per-layer-class timing of the factor SYRK (kfb_factor_update) at ResNet-50 shapes, batch 32.
Usage: python tools/bench_factor.py [mma_mode]   -> one line per class: ms, TFLOP/s, GB/s."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from distributed_kfac_pytorch_b200 import _native as N  # noqa: E402
from distributed_kfac_pytorch_b200.layers.base import alloc2d  # noqa: E402

DEV = 'cuda:0'
NB = 32
# (tag, kind, C, H(in), k, stride, pad, count in ResNet-50)
CLASSES = [
    ('A3x3_c256_14', 'A', 256, 14, 3, 1, 1, 5), ('A3x3_c64_56', 'A', 64, 56, 3, 1, 1, 3),
    ('A3x3_c128_28', 'A', 128, 28, 3, 1, 1, 3), ('A3x3_c512_7', 'A', 512, 7, 3, 1, 1, 2),
    ('A3x3s2_c128_56', 'A', 128, 56, 3, 2, 1, 1), ('A3x3s2_c256_28', 'A', 256, 28, 3, 2, 1, 1),
    ('A3x3s2_c512_14', 'A', 512, 14, 3, 2, 1, 1), ('A7x7_conv1', 'A', 3, 224, 7, 2, 3, 1),
    ('G_c1024_14', 'G', 1024, 14, 1, 1, 0, 7), ('A1x1_c1024_14', 'G', 1024, 14, 1, 1, 0, 6),
    ('G_c512_28', 'G', 512, 28, 1, 1, 0, 9), ('G_c256_56', 'G', 256, 56, 1, 1, 0, 7),
    ('G_c2048_7', 'G', 2048, 7, 1, 1, 0, 6), ('G_c64_112', 'G', 64, 112, 1, 1, 0, 1),
    ('G_c64_56', 'G', 64, 56, 1, 1, 0, 11), ('A1x1s2_c256_56', 'A', 256, 56, 1, 2, 0, 1),
]


def main():
    mode = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    ops = N.get_ops()
    ops._enter(torch.zeros(1, device=DEV))
    ops.set_mma_mode(mode)
    only = sys.argv[2] if len(sys.argv) > 2 else ''
    tot_ms = 0.0
    for tag, kind, c, h, k, s, p, count in CLASSES:
        if only and only not in tag:
            continue
        x = torch.randn(NB, c, h, h, device=DEV)
        if kind == 'A':
            oh = (h + 2 * p - k) // s + 1
            S = oh * oh
            seg = N.Segment(N.KFB_SEG_IM2COL, x, NB, c, 1.0 / (S * S * NB * S), h=h, w=h, kernel=(k, k),
                            stride=(s, s), padding=(p, p))
            d = c * k * k
        else:
            S = h * h
            seg = N.Segment(N.KFB_SEG_NCHW, x, NB, c, 1.0 / (S * S * NB * S), h=h, w=h)
            d = c
        K = NB * S
        _, st = alloc2d(d, d, DEV, zero=True)
        job = [N.FactorJob(d, [seg], st, False, True, None)]
        for _ in range(2):
            ops.factor_update(job, 0.95)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            ops.factor_update(job, 0.95)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        flops = K * d * (d + 1)
        gb = x.numel() * 4 + 3 * d * d * 4
        tot_ms += ms * count
        print('%-16s d=%5d K=%7d  %8.3f ms  %7.1f TFLOP/s  %7.1f GB/s   x%d' %
              (tag, d, K, ms, flops / ms / 1e9, gb / ms / 1e6, count))
    print('weighted total (sum ms x count): %.2f ms' % tot_ms)


if __name__ == '__main__':
    main()
