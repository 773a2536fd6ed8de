"""Host logic of the drop-in package on CPU: the native ops are replaced by the
oracle-backed stand-in (tests/_cpu_ops.py), everything else is the product code.
Checked against fixtures produced by the unmodified reference (tests/golden)."""
import json
import os
import warnings

import pytest
import torch
import torch.nn.functional as F

from conftest import GOLDEN, load_golden
import models
import distributed_kfac_pytorch_b200 as kfac
from distributed_kfac_pytorch_b200 import _native, comm, utils
from _cpu_ops import OracleOps

GRAD_TOL = dict(rtol=1e-2, atol=1e-4)      # north_star: preconditioned gradients
FACTOR_TOL = dict(rtol=1e-3, atol=1e-6)    # north_star: factors


@pytest.fixture()
def cpu_ops():
    ops = OracleOps()
    _native.set_ops(ops)
    comm.reset_comm_backend()
    yield ops
    _native.set_ops(None)
    comm.reset_comm_backend()


def _run_like_golden(rec, model, batch_fn, loss_fn, kfac_kwargs):
    model.load_state_dict(rec['init_state'])
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    pre = kfac.KFAC(model, **kfac_kwargs)
    for s, gold in enumerate(rec['steps']):
        x, y = batch_fn(s)
        opt.zero_grad()
        loss = loss_fn(model(x), y)
        loss.backward()
        assert abs(loss.item() - gold['loss']) < 1e-3 * max(1.0, abs(gold['loss']))
        pre.step()
        for i, layer in enumerate(pre.layers):
            torch.testing.assert_close(layer.state['A'], gold['A'][i], **FACTOR_TOL)
            torch.testing.assert_close(layer.state['G'], gold['G'][i], **FACTOR_TOL)
            w, b = gold['grads_out'][i]
            torch.testing.assert_close(layer.module.weight.grad, w, **GRAD_TOL)
            if b is not None:
                torch.testing.assert_close(layer.module.bias.grad, b, **GRAD_TOL)
        opt.step()
    return pre


def test_cnn_matches_reference(cpu_ops):
    rec = load_golden('cnn_single.pt')
    pre = _run_like_golden(rec, models.make_cnn(), models.cnn_batch, F.cross_entropy,
                           rec['kfac_kwargs'])
    assert [repr(l) for l in pre.layers] == rec['layer_reprs']
    assert sorted(pre.state_dict().keys()) == rec['state_dict_keys']
    # one native call per stage, not per layer
    assert ('factor_update', 12) in cpu_ops.calls and ('precondition', 6) in cpu_ops.calls


def test_lstm_matches_reference(cpu_ops):
    rec = load_golden('lstm_single.pt')
    pre = _run_like_golden(rec, models.make_lstm(kfac.modules.LSTM), models.lstm_batch,
                           F.nll_loss, rec['kfac_kwargs'])
    assert len(pre.layers) == 4
    assert all(l.__class__.__name__ == 'LinearMultiLayer' for l in pre.layers)


def test_state_dict_roundtrip_and_layout(cpu_ops):
    rec = load_golden('cnn_single.pt')
    pre = _run_like_golden(rec, models.make_cnn(), models.cnn_batch, F.cross_entropy,
                           rec['kfac_kwargs'])
    sd = pre.state_dict()
    assert len(sd['layers']) == 6
    for entry, layer in zip(sd['layers'], pre.layers):
        assert set(entry.keys()) == {'A', 'G'}
        assert entry['A'].shape == (layer.state['A'].shape[0],) * 2
        assert entry['A'].dtype == torch.float32
        for t in entry.values():      # saved tensors are plain contiguous d x d copies, not padded views
            assert t.is_contiguous() and t.untyped_storage().nbytes() == t.numel() * 4
    full = pre.state_dict(include_layer_inverses=True)['layers'][1]    # c2: da = 73, pitch-padded inside
    assert {'QA', 'QG', 'dGdA'} <= set(full.keys())
    assert all(t.is_contiguous() and t.untyped_storage().nbytes() == t.numel() * t.element_size()
               for t in full.values() if torch.is_tensor(t))
    # legacy key names are accepted (kfac/layers/base.py:64-68)
    legacy = dict(sd)
    legacy['layers'] = [{'A_factor': e['A'].clone(), 'G_factor': e['G'].clone()} for e in sd['layers']]
    model2 = models.make_cnn()
    pre2 = kfac.KFAC(model2, **rec['kfac_kwargs'])
    pre2.load_state_dict(legacy)
    assert pre2.param_groups[0]['step'] == pre.param_groups[0]['step']
    for a, b in zip(pre.layers, pre2.layers):
        torch.testing.assert_close(a.state['A'], b.state['A'], rtol=0, atol=0)
        assert 'QA' in b.state and 'dGdA' in b.state     # inverses recomputed on load
    with pytest.raises(ValueError):
        bad = dict(sd)
        bad['layers'] = sd['layers'][:-1]
        pre2.load_state_dict(bad)


def test_argument_validation(cpu_ops):
    m = models.make_cnn()
    for kw in (dict(lr=-1), dict(factor_decay=0), dict(factor_decay=1.5), dict(damping=0),
               dict(kl_clip=0), dict(factor_update_freq=0), dict(inv_update_freq=0),
               dict(assignment_strategy='x')):
        with pytest.raises(ValueError):
            kfac.KFAC(m, **kw)
    with pytest.warns(UserWarning):
        p = kfac.KFAC(m, comm_method=kfac.CommMethod.MEM_OPT)
    assert p.distribute_layer_factors is False and p.grad_worker_fraction == 0
    p = kfac.KFAC(models.make_cnn(), skip_layers='Linear')
    assert len(p.layers) == 5
    assert 'registered_layers: 5' in repr(p)
    with pytest.raises(RuntimeError):
        p.layers[0].update_gradient()
    with pytest.raises(ValueError):
        p.layers[0].compute_A_inv()


def test_eval_and_off_steps_do_not_capture(cpu_ops):
    m = models.make_cnn()
    p = kfac.KFAC(m, factor_update_freq=2, inv_update_freq=2)
    x, y = models.cnn_batch(0)
    with torch.no_grad():
        m(x)
    assert all(len(l.a_inputs) == 0 for l in p.layers)
    F.cross_entropy(m(x), y).backward()
    assert all(len(l.a_inputs) == 1 and len(l.g_outputs) == 1 for l in p.layers)
    p.step()
    assert all(len(l.a_inputs) == 0 for l in p.layers)
    m.zero_grad()
    F.cross_entropy(m(x), y).backward()      # step 1: not a factor step
    assert all(len(l.a_inputs) == 0 and len(l.g_outputs) == 0 for l in p.layers)


def test_scheduler(cpu_ops):
    p = kfac.KFAC(models.make_cnn(), damping=0.01, factor_update_freq=10, inv_update_freq=100)
    s = kfac.KFACParamScheduler(p, damping_alpha=0.5, damping_schedule=[2, 4],
                                update_freq_alpha=2, update_freq_schedule=[3])
    vals = []
    for _ in range(5):
        s.step()
        g = p.param_groups[0]
        vals.append((round(g['damping'], 6), g['factor_update_freq'], g['inv_update_freq']))
    assert vals == [(0.01, 10, 100), (0.005, 10, 100), (0.005, 20, 200), (0.0025, 20, 200),
                    (0.0025, 20, 200)]
    sd = s.state_dict()
    assert 'kfac' not in sd and not any('func' in k for k in sd)


def test_lr_scheduler_wraps_kfac(cpu_ops):
    p = kfac.KFAC(models.make_cnn(), lr=0.1)
    sched = torch.optim.lr_scheduler.LambdaLR(p, lambda e: 0.5 ** e)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        sched.step()
    assert abs(p.param_groups[0]['lr'] - 0.05) < 1e-12


# ---- placement: bit-exact -------------------------------------------------------

def test_utils_kats():
    from test_oracle_vs_golden import LOAD_BALANCE_KATS, GRAD_RANK_KATS, INV_RANK_KATS
    for n, work, expected in LOAD_BALANCE_KATS:
        assert utils.load_balance(n, work) == expected
    with pytest.raises(ValueError):
        utils.load_balance(1, [])
    for args, expected in GRAD_RANK_KATS:
        assert utils.partition_grad_ranks(*args) == expected
    for args, expected in INV_RANK_KATS:
        assert utils.partition_inv_ranks(*args) == expected
    assert utils.get_block_boundary(1, 2, [4, 6]) == ([2, 3], [4, 6])
    with pytest.raises(ValueError):
        utils.get_block_boundary(2, 2, [4, 4])


def test_placement_matches_reference_bit_exact():
    comm.reset_comm_backend()
    comm.init_comm_backend()
    with open(os.path.join(GOLDEN, 'allocator.json')) as f:
        gold = json.load(f)
    for name in ('resnet50', 'resnet152'):
        dims = gold['load_balance'][name + '/dims']
        for strategy, p in (('compute', 3), ('memory', 2)):
            for world in (2, 4, 8):
                costs = [a ** p + g ** p for a, g in dims]
                assert utils.load_balance(world, costs) == \
                    gold['load_balance']['%s/%s/W%d' % (name, strategy, world)]['locs']
    for key, ref in gold['groups'].items():
        world, frac = key.split('/')
        world, frac = int(world[1:]), float(frac[1:])
        if 'error' in ref:
            with pytest.raises(ValueError):
                utils.WorkerAllocator(world, frac)
            continue
        alloc = utils.WorkerAllocator(world, frac)
        assert alloc.bcast_grad_ranks == ref['grad_ranks']
        assert alloc.bcast_inv_ranks == ref['inv_ranks']
        for inv_rank, r in ref['per_inv_rank'].items():
            src = alloc.get_inv_ranks(int(inv_rank))
            assert src == r['inv_ranks']
            assert [[s, g.ranks] for s, g in alloc.get_grad_groups(src)] == r['grad_src_group']
    comm.reset_comm_backend()


# ---- two ranks over gloo ----------------------------------------------------------

def _world_worker(rank, world, port, method_name, fraction, extra, out_path):
    import torch.distributed as dist
    import world_replay
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), LOCAL_RANK=str(rank))
    torch.set_num_threads(1)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    _native.set_ops(OracleOps())
    comm.reset_comm_backend()
    rec = world_replay.replay(kfac, method_name, fraction, torch.device('cpu'), rank, extra)
    pre = rec.pop('pre')
    rec['vec_numel'] = [(l.state['dA'].numel(), l.state['dG'].numel())
                        if l.state.get('dA') is not None and l.state.get('dG') is not None else None
                        for l in pre.layers]
    rec['dims'] = [l._dims() for l in pre.layers]
    torch.save(rec, out_path % rank)
    dist.barrier()
    dist.destroy_process_group()


def _spawn_world(tmp_path, world, port, method, fraction, extra=None):
    import torch.multiprocessing as mp
    out = str(tmp_path / 'rank%d.pt')
    mp.spawn(_world_worker, args=(world, port, method, fraction, extra, out), nprocs=world, join=True)
    return [torch.load(out % r, weights_only=False) for r in range(world)]


@pytest.mark.parametrize('world', [2, 4])
@pytest.mark.parametrize('idx,method,fraction', [(0, 'COMM_OPT', 1.0), (1, 'MEM_OPT', 0.0),
                                                 (2, 'HYBRID_OPT', 0.5)])
def test_world_gloo_matches_reference(tmp_path, world, idx, method, fraction):
    import world_replay
    gold = world_replay.load_world_golden(world)[method]
    ranks = _spawn_world(tmp_path, world, 29711 + 10 * world + idx, method, fraction)
    for r in ranks:
        world_replay.compare(r, gold)     # assignment bit-exact, factors / gradients within tolerance
    # every rank ends with the same gradients
    for r in ranks[1:]:
        for i in range(6):
            torch.testing.assert_close(ranks[0]['steps'][2]['grads_out'][i][0],
                                       r['steps'][2]['grads_out'][i][0], rtol=1e-5, atol=1e-7)
    if method == 'MEM_OPT':
        # only the inverse rank holds a layer's eigenvectors (memory story of SURVEY §8c)
        for i, a in enumerate(gold['assign']):
            for rank in range(world):
                has = 'QA' in ranks[rank]['held'][i]
                assert has == (rank == a['a_rank'])


@pytest.mark.parametrize('idx,key,method,fraction', [(0, 'COMM_OPT/no_prediv', 'COMM_OPT', 1.0),
                                                     (1, 'HYBRID_OPT/no_prediv', 'HYBRID_OPT', 0.5),
                                                     (2, 'COMM_OPT/distributed', 'COMM_OPT', 1.0),
                                                     (3, 'COMM_OPT/inv_fp16', 'COMM_OPT', 1.0)])
def test_world2_switches_match_reference(tmp_path, idx, key, method, fraction):
    """Eigenvalues travelling instead of dGdA (precompute_outer_eigen=False: the 1-D entries of the
    fused inverse slab), A / G decomposed on different ranks (distribute_layer_factors=True,
    kfac/preconditioner.py:638-645), and inverses that travel in 16 bits (inv_dtype=float16)."""
    import world_replay
    gold = load_golden('cnn_world2_variants.pt')[key]
    ranks = _spawn_world(tmp_path, 2, 29761 + idx, method, fraction, gold['kwargs'])
    for r in ranks:
        world_replay.compare(r, gold)
        n_held = 0 if 'no_prediv' in key or 'distributed' in key else 1
        for (da, dg), vec in zip(r['dims'], r['vec_numel']):
            if vec is not None:           # this rank is a gradient worker of the layer
                assert vec == (da, dg)    # dA / dG views carved out of the slab have their full length
                n_held += 1
        assert n_held > 0


# ---- SURVEY §8f rank 2/3 switches: in-hook factors, accumulated micro-batches, no prediv -------

def _run_variant(name, device='cpu'):
    rec = load_golden('cnn_variants.pt')[name]
    model = models.make_cnn().to(device)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    kw = dict(damping=0.003, factor_decay=0.95, factor_update_freq=1, inv_update_freq=2, kl_clip=0.001,
              lr=0.1)
    scale = rec.get('scale', 1.0)
    extra = dict(rec['kwargs'])
    if scale != 1.0:
        extra['grad_scaler'] = models.FixedScaler(scale)
    pre = kfac.KFAC(model, **dict(kw, **extra))
    for s, gold in enumerate(rec['steps']):
        opt.zero_grad()
        for m in range(rec['micro']):
            x, y = models.cnn_batch(10 * s + m)
            (F.cross_entropy(model(x.to(device)), y.to(device)) * scale / rec['micro']).backward()
        if scale != 1.0:
            for p in model.parameters():
                p.grad.div_(scale)
        pre.step()
        for i, layer in enumerate(pre.layers):
            torch.testing.assert_close(layer.state['A'].cpu(), gold['A'][i], **FACTOR_TOL)
            torch.testing.assert_close(layer.state['G'].cpu(), gold['G'][i], **FACTOR_TOL)
            torch.testing.assert_close(layer.module.weight.grad.cpu(), gold['grads_out'][i][0], **GRAD_TOL)
        opt.step()


@pytest.mark.parametrize('name', ['in_hook', 'accumulate', 'no_prediv', 'grad_scaler', 'inverse'])
def test_variants_match_reference(cpu_ops, name):
    _run_variant(name)


# ---- 16-bit inverses, GradScaler captures that cannot be unscaled -----------------

def _run_amp_variant(name, device='cpu', scaler=None):
    rec = load_golden('cnn_amp.pt')[name]
    model = models.make_cnn().to(device)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    kw = dict(damping=0.003, factor_decay=0.95, factor_update_freq=1, inv_update_freq=2, kl_clip=0.001,
              lr=0.1)
    scale = rec['scale']
    extra = dict(rec['kwargs'])
    if scale != 1.0:
        extra['grad_scaler'] = scaler if scaler is not None else models.FixedScaler(scale)
    pre = kfac.KFAC(model, **dict(kw, **extra))
    for s, gold in enumerate(rec['steps']):
        opt.zero_grad()
        for m in range(rec['micro']):
            x, y = models.cnn_batch(10 * s + m)
            (F.cross_entropy(model(x.to(device)), y.to(device)) * scale / rec['micro']).backward()
        if scale != 1.0:
            for p in model.parameters():
                p.grad.div_(scale)
        if rec.get('poison_step') == s:
            models.poison_captures(pre.layers)
        pre.step()
        for i, layer in enumerate(pre.layers):
            torch.testing.assert_close(layer.state['A'].cpu(), gold['A'][i], **FACTOR_TOL)
            torch.testing.assert_close(layer.state['G'].cpu(), gold['G'][i], **FACTOR_TOL)
            torch.testing.assert_close(layer.module.weight.grad.cpu(), gold['grads_out'][i][0].float(), **GRAD_TOL)
        opt.step()
    return pre


def test_inv_dtype_float16_matches_reference(cpu_ops):
    pre = _run_amp_variant('inv_fp16')
    q = pre.layers[1].state['QA']
    assert torch.equal(q, q.half().float())          # what is stored are 16-bit values
    with pytest.raises(NotImplementedError):
        kfac.KFAC(models.make_cnn(), inv_dtype=torch.float64)


def test_unscalable_captures_are_dropped_like_the_reference(cpu_ops):
    with pytest.warns(UserWarning, match='discarded'):
        _run_amp_variant('grad_scaler_inf')



# ---- register_shared_module (kfac/preconditioner.py:404-470) --------------------------------------

def _run_shared(device='cpu'):
    rec = load_golden('shared_module.pt')
    model = models.make_tied().to(device)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    pre = kfac.KFAC(model, **rec['kwargs'])
    with pytest.warns(UserWarning):
        pre.register_shared_module(model.first, model.second)
    assert len(pre.layers) == rec['n_layers'] == 2
    for s, gold in enumerate(rec['steps']):
        x, y = models.tied_batch(s)
        opt.zero_grad()
        F.cross_entropy(model(x.to(device)), y.to(device)).backward()
        pre.step()
        for i, layer in enumerate(pre.layers):
            torch.testing.assert_close(layer.state['A'].cpu(), gold['A'][i], **FACTOR_TOL)
            torch.testing.assert_close(layer.state['G'].cpu(), gold['G'][i], **FACTOR_TOL)
            torch.testing.assert_close(layer.module.weight.grad.cpu(), gold['grads_out'][i][0], **GRAD_TOL)
            torch.testing.assert_close(layer.module.bias.grad.cpu(), gold['grads_out'][i][1], **GRAD_TOL)
        opt.step()
    return pre


def test_shared_module_matches_reference(cpu_ops):
    pre = _run_shared()
    with pytest.raises(ValueError):      # needs accumulate_data
        kfac.KFAC(models.make_tied()).register_shared_module(torch.nn.Linear(2, 2), torch.nn.Linear(2, 2))
    with pytest.raises(ValueError):
        pre.register_shared_module(object(), torch.nn.Linear(2, 2))
