"""Host-side cost of KFAC.step() on ResNet-50 (time until the call returns, nothing synchronised)
next to its device time, for ordinary steps and for the factor (+eig) step, and of the two big
C-ABI calls alone."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402


def main():
    import distributed_kfac_pytorch_b200 as kfac
    from distributed_kfac_pytorch_b200 import _native
    device = torch.device('cuda:0')
    torch.cuda.set_device(device)
    ops = _native.get_ops()
    ops._enter(torch.zeros(1, device=device))
    model = bench.build_model('resnet50').to(device)
    pre = kfac.KFAC(model, **dict(bench.KFAC_KW, **bench.comm_kwargs('comm-opt')))
    x, y = bench.synthetic_batch(0, device)
    caps, grads = bench.capture(model, pre, x, y)
    calls = {}
    for name in ('factor_update', 'precondition', 'syevd_batch', 'scale_writeback', 'tri_pack', 'tri_unpack'):
        inner = getattr(ops, name)

        def wrap(*a, _inner=inner, _name=name, **kw):
            t0 = time.perf_counter()
            out = _inner(*a, **kw)
            calls.setdefault(_name, []).append((time.perf_counter() - t0) * 1e3)
            return out
        setattr(ops, name, wrap)
    rows = []
    for step in range(23):
        bench.feed(pre, caps)
        bench.restore_grads(pre, grads)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        pre.step()
        e1.record()
        host = (time.perf_counter() - t0) * 1e3
        torch.cuda.synchronize()
        rows.append((step, host, e0.elapsed_time(e1)))
    for step, host, dev in rows:
        if step >= 10:
            print('step %2d  host %8.3f ms   device window %8.3f ms' % (step, host, dev))
    for name, v in calls.items():
        v = v[len(v) // 2:]
        print('%-16s host time per call: mean %.3f ms over %d calls' % (name, sum(v) / len(v), len(v)))


if __name__ == '__main__':
    main()
