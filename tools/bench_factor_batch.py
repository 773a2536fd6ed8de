"""Time the whole factor stage (all 108 ResNet-50 factors in ONE kfb_factor_update call: staging copies,
SYRK, running average) back to back, GPU busy: python tools/bench_factor_batch.py [model] [reps] [fp16|bf16]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import distributed_kfac_pytorch_b200 as kfac  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else 'resnet50'
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    half = {'fp16': torch.float16, 'bf16': torch.bfloat16}.get(sys.argv[3] if len(sys.argv) > 3 else '')
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(0)
    model = bench.build_model(name).to(dev)
    pre = kfac.KFAC(model, **dict(bench.KFAC_KW, **bench.model_kwargs(name)))
    x, y = bench.synthetic_batch(0, dev, name)
    caps, grads = bench.capture(model, pre, x, y)
    if half is not None:      # what the hooks capture under autocast: 16-bit activations and output gradients
        caps = [([t.to(half) for t in a], [t.to(half) for t in g]) for a, g in caps]
    work = None
    for _ in range(3):
        bench.feed(pre, caps)
        pre.compute_factors(alpha=0.95)
    torch.cuda.synchronize()
    work = bench.algorithmic_work(pre, caps)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for l, (a, g) in zip(pre.layers, caps):
            l.a_inputs, l.g_outputs = list(a), list(g)
        pre.compute_factors(alpha=0.95)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print('%s%s: factor stage %.3f ms per call, %.1f TFLOP/s (upper-triangle FLOPs %.1f G)'
          % (name, '' if half is None else ' (%s captures)' % sys.argv[3], ms, work['syrk_flops'] / ms / 1e9, work['syrk_flops'] / 1e9))


if __name__ == '__main__':
    main()
