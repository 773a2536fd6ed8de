"""Two ranks over NCCL on two GPUs (skipped on a one-GPU box): all three CommMethods reproduce
the gradients the unmodified reference produced on two gloo ranks (tests/golden/cnn_world2.pt),
and the worker-to-layer assignment is bit-exact."""
import os
import warnings

import pytest
import torch
import torch.nn.functional as F

from conftest import load_golden
import models

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, method_name, fraction, out_path):
    import torch.distributed as dist
    import distributed_kfac_pytorch_b200 as kfac
    from distributed_kfac_pytorch_b200 import _native, comm
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    _native.set_ops(None)
    comm.reset_comm_backend()
    model = models.make_cnn().to(dev)
    ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[rank])
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        pre = kfac.KFAC(model, comm_method=getattr(kfac.CommMethod, method_name),
                        grad_worker_fraction=fraction, distribute_layer_factors=False,
                        damping=0.003, factor_decay=0.95, factor_update_freq=1,
                        inv_update_freq=1, kl_clip=0.001, lr=0.1)
    steps = []
    for s in range(3):
        x, y = models.cnn_batch(s, rank)
        opt.zero_grad()
        F.cross_entropy(ddp(x.to(dev)), y.to(dev)).backward()
        pre.step()
        steps.append({'grads_out': [l.module.weight.grad.cpu().clone() for l in pre.layers],
                      'A': [l.state['A'].cpu().clone() for l in pre.layers]})
        opt.step()
    assign = [{'a_rank': l.compute_A_inv_rank, 'g_rank': l.compute_G_inv_rank,
               'grad_ranks': list(l.compute_grad_ranks),
               'grad_src_group': [(src, list(grp.ranks)) for src, grp in l.broadcast_grad_groups]}
              for l in pre.layers]
    torch.save({'steps': steps, 'assign': assign}, out_path % rank)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize('idx,method,fraction', [(0, 'COMM_OPT', 1.0), (1, 'MEM_OPT', 0.0),
                                                 (2, 'HYBRID_OPT', 0.5)])
def test_two_gpus_nccl_match_reference(tmp_path, idx, method, fraction):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    import torch.multiprocessing as mp
    gold = load_golden('cnn_world2.pt')[method]
    out = str(tmp_path / 'rank%d.pt')
    mp.spawn(_worker, args=(2, 29811 + idx, method, fraction, out), nprocs=2, join=True)
    ranks = [torch.load(out % r, weights_only=False) for r in range(2)]
    for r in ranks:
        assert r['assign'] == gold['assign']
        for s in range(3):
            for i in range(6):
                torch.testing.assert_close(r['steps'][s]['A'][i], gold['steps'][s]['A'][i], rtol=1e-3, atol=1e-6)
                torch.testing.assert_close(r['steps'][s]['grads_out'][i], gold['steps'][s]['grads_out'][i][0],
                                           rtol=1e-2, atol=1e-4)
    for i in range(6):
        torch.testing.assert_close(ranks[0]['steps'][2]['grads_out'][i], ranks[1]['steps'][2]['grads_out'][i],
                                   rtol=1e-5, atol=1e-7)
