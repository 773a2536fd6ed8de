"""Measure the dense TF32 tensor-core peak of this GPU the way MEASURED_PEAKS.json measures bf16:
torch.matmul (cuBLAS) on 8192^3 fp32 operands with allow_tf32, best of 10 (burst) and back to back
for 4 s (sustained), SM clock sampled during the sustained part.  Prints one JSON object; bench.py
reads the committed copy (profiles/r2_tf32_peak.json) for its roofline denominators."""
import json
import subprocess
import sys
import time

import torch


def sm_clock():
    try:
        out = subprocess.run(['nvidia-smi', '-i', '0', '--query-gpu=clocks.sm,clocks.max.sm,power.draw',
                              '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=10).stdout
        return [float(x) for x in out.strip().split(',')]
    except Exception:
        return None


def main():
    n = 8192
    torch.backends.cuda.matmul.allow_tf32 = True
    a = torch.randn(n, n, device='cuda')
    b = torch.randn(n, n, device='cuda')
    c = torch.empty(n, n, device='cuda')
    flops = 2.0 * n ** 3
    for _ in range(3):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
        time.sleep(0.05)
    burst = flops / (best * 1e-3) / 1e12
    clocks = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    count = 0
    e0.record()
    while time.time() - t0 < 4.0:
        for _ in range(50):
            torch.matmul(a, b, out=c)
        count += 50
        if count % 500 == 0:
            clocks.append(sm_clock())
    e1.record()
    torch.cuda.synchronize()
    sustained = flops * count / (e0.elapsed_time(e1) * 1e-3) / 1e12
    # the same measurement for bf16, to tie this run to MEASURED_PEAKS.json
    ab, bb = a.bfloat16(), b.bfloat16()
    cb = torch.empty(n, n, device='cuda', dtype=torch.bfloat16)
    for _ in range(3):
        torch.matmul(ab, bb, out=cb)
    torch.cuda.synchronize()
    bestb = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(ab, bb, out=cb)
        e1.record()
        torch.cuda.synchronize()
        bestb = min(bestb, e0.elapsed_time(e1))
        time.sleep(0.05)
    clocks = [c for c in clocks if c]
    print(json.dumps({'tf32_tflops': burst, 'tf32_tflops_sustained': sustained,
                      'bf16_tflops_same_run': flops / (bestb * 1e-3) / 1e12,
                      'how': 'torch.matmul fp32 operands, allow_tf32=True, 8192^3 (2*N^3): best of 10 (burst) and back '
                             'to back for 4 s (sustained), CUDA events',
                      'sm_mhz_sustained': sorted(c[0] for c in clocks)[len(clocks) // 2] if clocks else None,
                      'sm_max_mhz': clocks[0][1] if clocks else None,
                      'gpu_name': torch.cuda.get_device_name(0), 'torch': torch.__version__}))


if __name__ == '__main__':
    sys.exit(main())
