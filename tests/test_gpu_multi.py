"""Several ranks over NCCL, one GPU each (skipped where the box has fewer GPUs): all three
CommMethods reproduce the factors and gradients the unmodified reference produced on as many gloo
ranks (tests/golden/cnn_world{2,4,8}.pt), the worker-to-layer assignment is bit-exact, and the
multi-rank-only constructor switches match their goldens.  bench.py runs the same replay on the
driver's own torchrun launch ("parity_world")."""
import os

import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, method_name, fraction, extra, out_path):
    import torch.distributed as dist
    import distributed_kfac_pytorch_b200 as kfac
    from distributed_kfac_pytorch_b200 import _native, comm
    import world_replay
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    _native.set_ops(None)
    comm.reset_comm_backend()
    rec = world_replay.replay(kfac, method_name, fraction, dev, rank, extra)
    rec.pop('pre')
    torch.save(rec, out_path % rank)
    dist.barrier()
    dist.destroy_process_group()


def _spawn(tmp_path, world, port, method, fraction, extra=None):
    if torch.cuda.device_count() < world:
        pytest.skip('needs %d GPUs' % world)
    import torch.multiprocessing as mp
    out = str(tmp_path / 'rank%d.pt')
    mp.spawn(_worker, args=(world, port, method, fraction, extra, out), nprocs=world, join=True)
    return [torch.load(out % r, weights_only=False) for r in range(world)]


@pytest.mark.parametrize('world', [2, 4, 8])
@pytest.mark.parametrize('idx,method,fraction', [(0, 'COMM_OPT', 1.0), (1, 'MEM_OPT', 0.0),
                                                 (2, 'HYBRID_OPT', 0.5)])
def test_nccl_ranks_match_reference(tmp_path, world, idx, method, fraction):
    import world_replay
    gold = world_replay.load_world_golden(world)[method]
    ranks = _spawn(tmp_path, world, 29811 + 10 * world + idx, method, fraction)
    for r in ranks:
        world_replay.compare(r, gold)
    for r in ranks[1:]:
        for i in range(6):
            torch.testing.assert_close(ranks[0]['steps'][2]['grads_out'][i][0], r['steps'][2]['grads_out'][i][0],
                                       rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize('idx,key,method,fraction', [(0, 'COMM_OPT/no_prediv', 'COMM_OPT', 1.0),
                                                     (1, 'HYBRID_OPT/no_prediv', 'HYBRID_OPT', 0.5),
                                                     (2, 'COMM_OPT/distributed', 'COMM_OPT', 1.0),
                                                     (3, 'COMM_OPT/inv_fp16', 'COMM_OPT', 1.0)])
def test_two_gpus_switches_match_reference(tmp_path, idx, key, method, fraction):
    import world_replay
    gold = load_golden('cnn_world2_variants.pt')[key]
    for r in _spawn(tmp_path, 2, 29871 + idx, method, fraction, gold['kwargs']):
        world_replay.compare(r, gold)
