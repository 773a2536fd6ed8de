"""Replay of the multi-rank SmallCNN golden runs (tests/golden/cnn_world*.pt, written by the
UNMODIFIED reference on gloo ranks: tests/golden/make_golden.py) with this repository's KFAC.

Shared by the gloo tests on CPU (tests/test_host_logic.py, native ops replaced by the oracle
stand-in), the NCCL tests (tests/test_gpu_multi.py) and bench.py's ``parity_world`` check, which
runs it on the N GPUs of the driver's own torchrun launch.  Needs an initialised process group.
"""
import os
import sys
import warnings

import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, 'golden')
for _p in (HERE, GOLDEN):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import models  # noqa: E402

GRAD_TOL = dict(rtol=1e-2, atol=1e-4)      # north_star: preconditioned gradients
FACTOR_TOL = dict(rtol=1e-3, atol=1e-6)    # north_star: factors

METHODS = (('COMM_OPT', 1.0), ('MEM_OPT', 0.0), ('HYBRID_OPT', 0.5))
KFAC_KW = dict(damping=0.003, factor_decay=0.95, factor_update_freq=1, inv_update_freq=1,
               kl_clip=0.001, lr=0.1, distribute_layer_factors=False)


def load_world_golden(world):
    gold = torch.load(os.path.join(GOLDEN, 'cnn_world%d.pt' % world), weights_only=False)
    for name in gold:
        if gold[name].get('steps') is None:      # stored once: the reference's methods agree bit for bit
            gold[name]['steps'] = gold['COMM_OPT']['steps']
    return gold


def replay(kfac, method_name, fraction, device, rank, extra=None, steps=3):
    """Three training steps of SmallCNN under DDP + KFAC on this rank; returns what the goldens hold."""
    model = models.make_cnn().to(device)
    ddp = torch.nn.parallel.DistributedDataParallel(
        model, device_ids=[device.index] if device.type == 'cuda' else None)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    kw = dict(KFAC_KW)
    kw.update(extra or {})
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        pre = kfac.KFAC(model, comm_method=getattr(kfac.CommMethod, method_name),
                        grad_worker_fraction=fraction, **kw)
    out = []
    for s in range(steps):
        x, y = models.cnn_batch(s, rank)
        opt.zero_grad()
        F.cross_entropy(ddp(x.to(device)), y.to(device)).backward()
        pre.step()
        out.append({'grads_out': [(l.module.weight.grad.detach().cpu().clone(),
                                   None if l.module.bias is None else l.module.bias.grad.detach().cpu().clone())
                                  for l in pre.layers],
                    'A': [l.state['A'].cpu().clone() for l in pre.layers],
                    'G': [l.state['G'].cpu().clone() for l in pre.layers]})
        opt.step()
    assign = [{'a_rank': l.compute_A_inv_rank, 'g_rank': l.compute_G_inv_rank,
               'grad_ranks': list(l.compute_grad_ranks),
               'grad_src_group': [(src, list(grp.ranks)) for src, grp in l.broadcast_grad_groups]}
              for l in pre.layers]
    held = [sorted(k for k, v in l.state.items() if v is not None) for l in pre.layers]
    return {'steps': out, 'assign': assign, 'held': held, 'pre': pre}


def _norm_assign(assign):
    return [{'a_rank': a['a_rank'], 'g_rank': a['g_rank'], 'grad_ranks': list(a['grad_ranks']),
             'grad_src_group': [(s, list(g)) for s, g in a['grad_src_group']]} for a in assign]


def compare(rec, gold):
    """Raises AssertionError unless the assignment is bit-exact and factors / gradients are within
    north_star's tolerances of the reference's.  Returns (max |dgrad|, max |dfactor|)."""
    assert _norm_assign(rec['assign']) == _norm_assign(gold['assign']), 'worker assignment differs'
    worst_g = worst_f = 0.0
    for got, ref in zip(rec['steps'], gold['steps']):
        for i in range(len(ref['A'])):
            for key in ('A', 'G'):
                torch.testing.assert_close(got[key][i], ref[key][i], **FACTOR_TOL)
                worst_f = max(worst_f, (got[key][i] - ref[key][i]).abs().max().item())
            for j in (0, 1):
                if ref['grads_out'][i][j] is None:
                    continue
                torch.testing.assert_close(got['grads_out'][i][j], ref['grads_out'][i][j], **GRAD_TOL)
                worst_g = max(worst_g, (got['grads_out'][i][j] - ref['grads_out'][i][j]).abs().max().item())
    return worst_g, worst_f
