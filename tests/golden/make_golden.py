"""Generate golden fixtures by running the UNMODIFIED reference K-FAC.

Run in the build container only (the GPU box has no /root/reference):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

It puts /root/reference first on sys.path so ``import kfac`` is the reference,
installs the ``torch.symeig`` shim SURVEY §8(c) documents (torch >= 2.0 removed
the API used at kfac/layers/utils.py:63), and writes

    tests/golden/cnn_single.pt   single-process SmallCNN, 4 steps
    tests/golden/lstm_single.pt  single-process LSTM (LinearMultiLayer), 2 steps
    tests/golden/cnn_world2.pt   2-rank gloo, COMM_OPT / MEM_OPT / HYBRID_OPT
    tests/golden/allocator.json  worker assignments for ResNet-50/152, W = 2, 4, 8
    tests/golden/cnn_variants.pt  SmallCNN with compute_factor_in_hook / accumulate_data / no prediv
    tests/golden/resnet32_single.pt  BASELINE configs[0]: CIFAR ResNet-32, batch 128, 2 steps
    tests/golden/cnn_world4.pt, cnn_world8.pt   4- and 8-rank gloo, the three CommMethods (rank 0's view)
    tests/golden/cnn_world2_variants.pt   2-rank gloo: precompute_outer_eigen=False under COMM_OPT and
                                 HYBRID_OPT, and distribute_layer_factors=True (A and G on different ranks)
    tests/golden/shared_module.pt   register_shared_module (tied encoder/decoder weight), single process
    tests/golden/cnn_amp.pt      inv_dtype=float16, and GradScaler captures with injected inf / nan

Nothing here is imported by the product or by the tests; the tests only read the
files it wrote.
"""
import hashlib
import json
import os
import sys
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get('KFAC_REFERENCE', '/root/reference')
sys.path.insert(0, REF)
sys.path.insert(1, HERE)
sys.dont_write_bytecode = True
warnings.filterwarnings('ignore')

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402


def _symeig(t, eigenvectors=True, upper=True):
    d, q = torch.linalg.eigh(t, UPLO='U' if upper else 'L')
    return d, q.contiguous()


torch.symeig = _symeig

import kfac  # noqa: E402  (the reference)
import models  # noqa: E402

assert os.path.realpath(kfac.__file__).startswith(os.path.realpath(REF)), kfac.__file__

KFAC_ARGS = dict(damping=0.003, factor_decay=0.95, factor_update_freq=1, inv_update_freq=2,
                 kl_clip=0.001, lr=0.1)


def _clone_list(ts):
    return [t.detach().clone() for t in ts]


def _grads(layer):
    w = layer.module.weight.grad.detach().clone()
    b = layer.module.bias.grad.detach().clone() if layer.has_bias else None
    return (w, b)


def run_single(model, batch_fn, loss_fn, steps, kfac_kwargs):
    torch.set_num_threads(1)
    init_state = {k: v.clone() for k, v in model.state_dict().items()}
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    pre = kfac.KFAC(model, **kfac_kwargs)
    record = {'init_state': init_state, 'kfac_kwargs': {k: v for k, v in kfac_kwargs.items()},
              'steps': []}
    for s in range(steps):
        x, y = batch_fn(s)
        opt.zero_grad()
        loss = loss_fn(model(x), y)
        loss.backward()
        entry = {'loss': loss.item()}
        entry['a_inputs'] = [_clone_list(l.a_inputs) for l in pre.layers]
        entry['g_outputs'] = [_clone_list(l.g_outputs) for l in pre.layers]
        entry['grads_in'] = [_grads(l) for l in pre.layers]
        pre.step()
        entry['grads_out'] = [_grads(l) for l in pre.layers]
        entry['A'] = [l.state['A'].clone() for l in pre.layers]
        entry['G'] = [l.state['G'].clone() for l in pre.layers]
        record['steps'].append(entry)
        opt.step()
    record['layer_reprs'] = [repr(l) for l in pre.layers]
    record['state_dict_keys'] = sorted(pre.state_dict().keys())
    record['param_group'] = {k: v for k, v in pre.param_groups[0].items() if k != 'params'}
    return record


def gen_cnn_single():
    model = models.make_cnn()
    rec = run_single(model, models.cnn_batch, F.cross_entropy, 4, KFAC_ARGS)
    torch.save(rec, os.path.join(HERE, 'cnn_single.pt'))
    print('cnn_single: losses', [round(s['loss'], 4) for s in rec['steps']])


def gen_cnn_variants():
    """compute_factor_in_hook=True, and accumulate_data=True over two
    micro-batches per optimisation step (gradient accumulation, examples/cnn_utils/engine.py:33-65)."""
    torch.set_num_threads(1)
    out = {}
    for name, kw, micro in (('in_hook', dict(compute_factor_in_hook=True), 1),
                            ('accumulate', dict(accumulate_data=True), 2),
                            ('no_prediv', dict(precompute_outer_eigen=False), 1),
                            ('inverse', dict(use_eigen_decomp=False), 1),
                            ('grad_scaler', dict(grad_scaler=models.FixedScaler(1024.0)), 1)):
        model = models.make_cnn()
        opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
        pre = kfac.KFAC(model, **dict(KFAC_ARGS, **kw))
        scale = kw['grad_scaler'].get_scale() if 'grad_scaler' in kw else 1.0
        steps = []
        for s in range(3):
            opt.zero_grad()
            for m in range(micro):
                x, y = models.cnn_batch(10 * s + m)
                (F.cross_entropy(model(x), y) * scale / micro).backward()
            if scale != 1.0:       # what GradScaler.unscale_(optimizer) does before preconditioner.step()
                for p in model.parameters():
                    p.grad.div_(scale)
            pre.step()
            steps.append({'grads_out': [_grads(l) for l in pre.layers],
                          'A': [l.state['A'].clone() for l in pre.layers],
                          'G': [l.state['G'].clone() for l in pre.layers]})
            opt.step()
        out[name] = {'kwargs': {k: v for k, v in kw.items() if k != 'grad_scaler'}, 'micro': micro,
                     'scale': scale, 'steps': steps}
    torch.save(out, os.path.join(HERE, 'cnn_variants.pt'))
    print('cnn_variants: ok')


def gen_amp_variants():
    """inv_dtype=float16 (eigenvectors / dGdA stored, shipped and multiplied in 16
    bits: kfac/layers/base.py:435-441,459-470) and the GradScaler capture filter with injected inf / nan
    (base.py:397-407): one of two accumulated captures dropped for one layer, both dropped for another."""
    torch.set_num_threads(1)
    out = {}
    # -- inv_dtype = float16
    model = models.make_cnn()
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    pre = kfac.KFAC(model, **dict(KFAC_ARGS, inv_dtype=torch.float16))
    steps = []
    for s in range(3):
        opt.zero_grad()
        x, y = models.cnn_batch(10 * s)
        F.cross_entropy(model(x), y).backward()
        pre.step()
        steps.append({'grads_out': [_grads(l) for l in pre.layers],
                      'A': [l.state['A'].clone() for l in pre.layers],
                      'G': [l.state['G'].clone() for l in pre.layers]})
        opt.step()
    assert pre.layers[0].state['QA'].dtype == torch.float16
    out['inv_fp16'] = {'kwargs': dict(inv_dtype=torch.float16), 'micro': 1, 'scale': 1.0, 'steps': steps}
    # -- GradScaler with captures that cannot be unscaled
    scale = 1024.0
    model = models.make_cnn()
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    pre = kfac.KFAC(model, **dict(KFAC_ARGS, accumulate_data=True, grad_scaler=models.FixedScaler(scale)))
    steps = []
    for s in range(3):
        opt.zero_grad()
        for m in range(2):
            x, y = models.cnn_batch(10 * s + m)
            (F.cross_entropy(model(x), y) * scale / 2).backward()
        for p in model.parameters():
            p.grad.div_(scale)
        if s == 1:
            models.poison_captures(pre.layers)
        pre.step()
        steps.append({'grads_out': [_grads(l) for l in pre.layers],
                      'A': [l.state['A'].clone() for l in pre.layers],
                      'G': [l.state['G'].clone() for l in pre.layers]})
        opt.step()
    out['grad_scaler_inf'] = {'kwargs': dict(accumulate_data=True), 'micro': 2, 'scale': scale, 'steps': steps,
                              'poison_step': 1}
    torch.save(out, os.path.join(HERE, 'cnn_amp.pt'))
    print('cnn_amp: ok')


def gen_resnet32():
    """BASELINE configs[0]: ResNet-32, 32x32, batch 128, single process, COMM_OPT, CIFAR defaults
    (factor_update_freq 1, inv_update_freq 10, damping 0.003).  Only what the parity test needs
    is stored: the final gradients of each step and two factors, in fp16-safe fp32."""
    torch.set_num_threads(8)
    model = models.make_resnet32()
    # the weights are reproducible from the seed (make_resnet32); only a checksum is stored
    init_checksum = float(sum(p.double().abs().sum() for p in model.parameters()))
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    kw = dict(damping=0.003, factor_decay=0.95, factor_update_freq=1, inv_update_freq=10, kl_clip=0.001, lr=0.1)
    pre = kfac.KFAC(model, **kw)
    steps = []
    for s in range(2):
        x, y = models.cifar_batch(s)
        opt.zero_grad()
        loss = F.cross_entropy(model(x), y)
        loss.backward()
        pre.step()
        steps.append({'loss': loss.item(),
                      'grads_out': [l.module.weight.grad.detach().half() for l in pre.layers],   # fp16: 5e-4 relative
                      'bias_out': pre.layers[-1].module.bias.grad.detach().clone(),
                      'A_first': pre.layers[0].state['A'].clone(), 'G_first': pre.layers[0].state['G'].clone(),
                      'A_last': pre.layers[-1].state['A'].clone(), 'G_last': pre.layers[-1].state['G'].clone(),
                      'A_trace': [l.state['A'].diagonal().sum().item() for l in pre.layers],
                      'G_trace': [l.state['G'].diagonal().sum().item() for l in pre.layers]})
        opt.step()
    torch.save({'init_checksum': init_checksum, 'kfac_kwargs': kw, 'steps': steps, 'n_layers': len(pre.layers)},
               os.path.join(HERE, 'resnet32_single.pt'))
    print('resnet32: layers', len(pre.layers), 'losses', [round(s['loss'], 4) for s in steps])


def gen_lstm_single():
    model = models.make_lstm(kfac.modules.LSTM)
    kw = dict(KFAC_ARGS, inv_update_freq=1, batch_first=False, accumulate_data=True,
              skip_layers=['linear', 'embedding'])
    rec = run_single(model, models.lstm_batch, F.nll_loss, 2, kw)
    torch.save(rec, os.path.join(HERE, 'lstm_single.pt'))
    print('lstm_single: layers', len(rec['steps'][0]['A']),
          'losses', [round(s['loss'], 4) for s in rec['steps']])


def _world2_worker(rank, world, port, method_name, fraction, out, extra=None):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    os.environ['LOCAL_RANK'] = str(rank)
    torch.symeig = _symeig
    torch.set_num_threads(1)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    model = models.make_cnn()
    ddp = torch.nn.parallel.DistributedDataParallel(model)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    kw = dict(KFAC_ARGS, inv_update_freq=1, distribute_layer_factors=False)
    kw.update(extra or {})
    pre = kfac.KFAC(model, comm_method=getattr(kfac.CommMethod, method_name),
                    grad_worker_fraction=fraction, **kw)
    steps = []
    for s in range(3):
        x, y = models.cnn_batch(s, rank)
        opt.zero_grad()
        F.cross_entropy(ddp(x), y).backward()
        grads_in = [_grads(l) for l in pre.layers]
        pre.step()
        steps.append({'grads_in': grads_in, 'grads_out': [_grads(l) for l in pre.layers],
                      'A': [l.state['A'].clone() for l in pre.layers],
                      'G': [l.state['G'].clone() for l in pre.layers]})
        opt.step()
    assign = [{'a_rank': l.compute_A_inv_rank, 'g_rank': l.compute_G_inv_rank,
               'grad_ranks': list(l.compute_grad_ranks),
               'grad_src_group': [(src, list(grp.ranks)) for src, grp in l.broadcast_grad_groups]}
              for l in pre.layers]
    if rank == 0:
        torch.save({'steps': steps, 'assign': assign}, out)
    dist.barrier()
    dist.destroy_process_group()


def gen_cnn_world2():
    import torch.multiprocessing as mp
    result = {}
    for i, (name, frac) in enumerate([('COMM_OPT', 1.0), ('MEM_OPT', 0.0), ('HYBRID_OPT', 0.5)]):
        tmp = '/tmp/_golden_w2_%s.pt' % name
        mp.spawn(_world2_worker, args=(2, 29611 + i, name, frac, tmp), nprocs=2, join=True)
        result[name] = torch.load(tmp, weights_only=False)
        os.remove(tmp)
    # the reference yields identical gradients for all three methods
    for name in ('MEM_OPT', 'HYBRID_OPT'):
        for s in range(3):
            for (w0, b0), (w1, b1) in zip(result['COMM_OPT']['steps'][s]['grads_out'],
                                          result[name]['steps'][s]['grads_out']):
                assert torch.equal(w0, w1), name
    torch.save(result, os.path.join(HERE, 'cnn_world2.pt'))
    print('cnn_world2: ok, methods agree bit-for-bit')


def gen_cnn_world(world):
    """The three CommMethods on `world` gloo ranks; rank 0's factors and gradients (every rank ends
    with the same ones) and the assignment.  bench.py replays this on `world` GPUs over NCCL."""
    import torch.multiprocessing as mp
    result = {}
    for i, (name, frac) in enumerate([('COMM_OPT', 1.0), ('MEM_OPT', 0.0), ('HYBRID_OPT', 0.5)]):
        tmp = '/tmp/_golden_w%d_%s.pt' % (world, name)
        mp.spawn(_world2_worker, args=(world, 29631 + 10 * world + i, name, frac, tmp), nprocs=world, join=True)
        rec = torch.load(tmp, weights_only=False)
        os.remove(tmp)
        for st in rec['steps']:
            del st['grads_in']
        result[name] = rec
    for name in ('MEM_OPT', 'HYBRID_OPT'):
        for s in range(3):
            for (w0, b0), (w1, b1) in zip(result['COMM_OPT']['steps'][s]['grads_out'],
                                          result[name]['steps'][s]['grads_out']):
                assert torch.equal(w0, w1), name
            for key in ('A', 'G'):
                for f0, f1 in zip(result['COMM_OPT']['steps'][s][key], result[name]['steps'][s][key]):
                    assert torch.equal(f0, f1), name
        result[name]['steps'] = None        # bit-identical to COMM_OPT's: stored once
    torch.save(result, os.path.join(HERE, 'cnn_world%d.pt' % world))
    print('cnn_world%d: ok, methods agree bit-for-bit' % world)


def gen_cnn_world2_variants():
    """Constructor switches that only matter with more than one rank: eigenvalues
    shipped instead of dGdA (precompute_outer_eigen=False) and A / G decomposed on different ranks
    (distribute_layer_factors=True, kfac/preconditioner.py:638-645)."""
    import torch.multiprocessing as mp
    result = {}
    cases = [('COMM_OPT/no_prediv', 'COMM_OPT', 1.0, dict(precompute_outer_eigen=False)),
             ('HYBRID_OPT/no_prediv', 'HYBRID_OPT', 0.5, dict(precompute_outer_eigen=False)),
             ('COMM_OPT/distributed', 'COMM_OPT', 1.0,
              dict(precompute_outer_eigen=False, distribute_layer_factors=True)),
             ('COMM_OPT/inv_fp16', 'COMM_OPT', 1.0, dict(inv_dtype=torch.float16))]
    for i, (key, name, frac, extra) in enumerate(cases):
        tmp = '/tmp/_golden_w2v_%d.pt' % i
        mp.spawn(_world2_worker, args=(2, 29691 + i, name, frac, tmp, extra), nprocs=2, join=True)
        rec = torch.load(tmp, weights_only=False)
        os.remove(tmp)
        for st in rec['steps']:
            del st['grads_in']
        rec['kwargs'] = extra
        result[key] = rec
    assert any(a['a_rank'] != a['g_rank'] for a in result['COMM_OPT/distributed']['assign'])
    torch.save(result, os.path.join(HERE, 'cnn_world2_variants.pt'))
    print('cnn_world2_variants: ok')


def gen_shared_module():
    """register_shared_module (kfac/preconditioner.py:404-470): two Linear modules that share one
    weight, the second registered by hand."""
    torch.set_num_threads(1)
    model = models.make_tied()
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    kw = dict(KFAC_ARGS, inv_update_freq=1, accumulate_data=True)
    pre = kfac.KFAC(model, **kw)
    pre.register_shared_module(model.first, model.second)
    steps = []
    for s in range(3):
        x, y = models.tied_batch(s)
        opt.zero_grad()
        F.cross_entropy(model(x), y).backward()
        pre.step()
        steps.append({'grads_out': [_grads(l) for l in pre.layers],
                      'A': [l.state['A'].clone() for l in pre.layers],
                      'G': [l.state['G'].clone() for l in pre.layers]})
        opt.step()
    torch.save({'kwargs': kw, 'steps': steps, 'n_layers': len(pre.layers)}, os.path.join(HERE, 'shared_module.pt'))
    print('shared_module: layers', len(pre.layers))


def gen_allocator():
    from kfac import utils as kutils
    from kfac import comm as kcomm
    kcomm.init_comm_backend()
    out = {'load_balance': {}, 'groups': {}}
    for name in ('resnet50', 'resnet152'):
        dims = models.resnet_factor_dims(name)
        for strategy, p in (('compute', 3), ('memory', 2)):
            for world in (2, 4, 8):
                costs = [a ** p + g ** p for a, g in dims]
                locs = kutils.load_balance(world, costs)
                key = '%s/%s/W%d' % (name, strategy, world)
                out['load_balance'][key] = {
                    'locs': locs,
                    'sha256_16': hashlib.sha256(json.dumps(locs).encode()).hexdigest()[:16]}
                # distribute_layer_factors=True variant (a-costs then g-costs, preconditioner.py:638-641)
                costs2 = [a ** p for a, _ in dims] + [g ** p for _, g in dims]
                out['load_balance'][key + '/distributed'] = {'locs': kutils.load_balance(world, costs2)}
        out['load_balance'][name + '/dims'] = dims
    for world in (1, 2, 4, 8, 16):
        for frac in (1, 0, 0.5, 0.25, 0.125):
            try:
                alloc = kutils.WorkerAllocator(world, frac)
            except ValueError as e:
                out['groups']['W%d/f%s' % (world, frac)] = {'error': str(e)}
                continue
            entry = {'grad_ranks': alloc.bcast_grad_ranks, 'inv_ranks': alloc.bcast_inv_ranks,
                     'per_inv_rank': {}}
            for inv_rank in range(world):
                src = alloc.get_inv_ranks(inv_rank)
                pairs = alloc.get_grad_groups(src)
                entry['per_inv_rank'][str(inv_rank)] = {
                    'inv_ranks': src, 'grad_src_group': [(s, g.ranks) for s, g in pairs]}
            out['groups']['W%d/f%s' % (world, frac)] = entry
    with open(os.path.join(HERE, 'allocator.json'), 'w') as f:
        json.dump(out, f, indent=0, sort_keys=True)
    print('allocator: R50 W8 sha', out['load_balance']['resnet50/compute/W8']['sha256_16'])


if __name__ == '__main__':
    which = sys.argv[1:] or ['cnn', 'lstm', 'world2', 'alloc', 'resnet32', 'variants']
    if 'resnet32' in which:
        gen_resnet32()
    if 'variants' in which:
        gen_cnn_variants()
    if 'cnn' in which:
        gen_cnn_single()
    if 'lstm' in which:
        gen_lstm_single()
    if 'alloc' in which:
        gen_allocator()
    if 'world2' in which:
        gen_cnn_world2()
    if 'world4' in which:
        gen_cnn_world(4)
    if 'world8' in which:
        gen_cnn_world(8)
    if 'world2v' in which:
        gen_cnn_world2_variants()
    if 'shared' in which:
        gen_shared_module()
    if 'amp' in which:
        gen_amp_variants()
