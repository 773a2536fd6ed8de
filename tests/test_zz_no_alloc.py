"""kfb_factor_update_static — the factor update for hosts that may not allocate GPU memory.

Collected last on purpose (pytest orders files by name): if it fails, a run that stops at the
first failure has already been through every other test.

What is checked, through the C ABI on a real GPU:
  * two calls in a row on the same stream give the same running averages as the ordinary
    kfb_factor_update entry point given the same jobs (the second call starts from the first
    call's state and uses different inputs, so anything the first call left behind in the
    shared built-in scratch array would show up as an error);
  * the call makes no allocation through torch's allocator either (the ordinary entry point
    borrows its workspace from torch; this one must not);
  * jobs that need more scratch than the built-in array holds are refused with a message that
    names both sizes, and nothing is written.
"""
import pytest
import torch

from distributed_kfac_pytorch_b200 import _native as N
from distributed_kfac_pytorch_b200.layers.base import alloc2d

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _jobs(seed, states, seed_identity):
    """Three factors of different layout classes: im2col conv input with a bias column, NCHW
    output gradient, Linear rows.  Returns jobs writing into `states`."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(6, 5, 11, 9, generator=g).to(DEV)
    go = torch.randn(6, 40, 7, 7, generator=g).to(DEV)
    a = torch.randn(48, 130, generator=g).to(DEV)
    S1 = 6 * 5            # conv 3x3 stride 2 pad 1 on 11x9 -> 6x5
    segs = [
        (5 * 9 + 1, [N.Segment(N.KFB_SEG_IM2COL, x, 6, 5, 1.0 / (S1 * S1 * 6 * S1), has_bias=True, h=11, w=9,
                               kernel=(3, 3), stride=(2, 2), padding=(1, 1))]),
        (40, [N.Segment(N.KFB_SEG_NCHW, go, 6, 40, 1.0 / (49.0 * 49 * 6 * 49), h=7, w=7)]),
        (131, [N.Segment(N.KFB_SEG_ROWS, a, 48, 130, 1.0 / 48, has_bias=True, ld=130)]),
    ]
    tris = [torch.zeros(d * (d + 1) // 2, device=DEV) for d, _ in segs]
    jobs = [N.FactorJob(d, s, st, seed_identity, True, tri) for (d, s), st, tri in zip(segs, states, tris)]
    return jobs, tris


def _fresh_states():
    return [alloc2d(d, d, DEV, zero=True)[1] for d in (46, 40, 131)]


@pytest.mark.parametrize('mode', [0, 1], ids=['simt', 'tcgen05'])
def test_static_scratch_matches_workspace_entry_point(mode):
    N.set_ops(None)
    ops = N.get_ops()
    ops._enter(torch.zeros(1, device=DEV))
    ops.set_mma_mode(mode)
    try:
        ref_states, got_states = _fresh_states(), _fresh_states()
        for call, (seed, alpha) in enumerate([(1, 0.95), (2, 0.5)]):
            ref_jobs, ref_tris = _jobs(seed, ref_states, call == 0)
            got_jobs, got_tris = _jobs(seed, got_states, call == 0)
            ops.factor_update(ref_jobs, alpha)                      # caller-workspace entry point
            torch.cuda.synchronize()
            before = torch.cuda.memory_stats(DEV)['allocation.all.allocated']
            ops.factor_update_static(got_jobs, alpha)               # built-in scratch, no allocation
            after = torch.cuda.memory_stats(DEV)['allocation.all.allocated']
            torch.cuda.synchronize()
            assert after == before, 'the no-alloc entry point went through the torch allocator'
            for r, g_, rt, gt in zip(ref_states, got_states, ref_tris, got_tris):
                # same kernels, same inputs: only the order of the split-K float additions may differ
                scale = r.abs().max().item()
                torch.testing.assert_close(g_, r, rtol=1e-5, atol=1e-6 * scale)
                torch.testing.assert_close(gt, rt, rtol=1e-5, atol=1e-6 * scale)
                assert torch.equal(g_, g_.t())
    finally:
        ops.set_mma_mode(1)


def test_static_scratch_rejects_oversized_jobs():
    N.set_ops(None)
    ops = N.get_ops()
    ops._enter(torch.zeros(1, device=DEV))
    limit = ops.static_scratch_bytes()
    assert limit == 16 << 20
    d = 2304                                  # 2304^2 fp32 split-K accumulators = 21 MB > 16 MiB
    rows = 5000                               # more rows than one work item takes: the job is split along K
    x = torch.randn(rows, d, device=DEV)
    _, st = alloc2d(d, d, DEV, zero=True)
    job = N.FactorJob(d, [N.Segment(N.KFB_SEG_ROWS, x, rows, d, 1.0 / rows, ld=d)], st, True, True, None)
    with pytest.raises(N.NativeError, match=r'need \d+ bytes of scratch but the built-in array holds %d' % limit):
        ops.factor_update_static([job], 0.9)
    torch.cuda.synchronize()
    assert float(st.abs().sum()) == 0.0       # refused before anything ran


def test_static_scratch_drops_staging_copies_that_do_not_fit():
    """The tensor-core SYRK normally reads im2col captures through fp32 staging copies kept in the
    workspace (about 3x the capture for a 3x3 conv).  They are an optimisation: when they do not fit
    the built-in array the no-alloc entry point plans the call without them instead of refusing it."""
    N.set_ops(None)
    ops = N.get_ops()
    ops._enter(torch.zeros(1, device=DEV))
    ops.set_mma_mode(1)
    x = torch.randn(8, 64, 56, 56, generator=torch.Generator().manual_seed(3)).to(DEV)   # 6.4 MB capture
    S = 56 * 56
    d = 64 * 9

    def job(state):
        seg = N.Segment(N.KFB_SEG_IM2COL, x, 8, 64, 1.0 / (S * S * 8.0 * S), h=56, w=56, kernel=(3, 3), stride=(1, 1),
                        padding=(1, 1))
        return N.FactorJob(d, [seg], state, True, True, None)

    _, ref = alloc2d(d, d, DEV, zero=True)
    _, got = alloc2d(d, d, DEV, zero=True)
    ops.factor_update([job(ref)], 0.95)
    need = ops.lib.kfb_factor_workspace_bytes(ops._factor_cjobs([job(ref)])[0], 1)
    assert need > ops.static_scratch_bytes()          # with staging it would not fit
    ops.factor_update_static([job(got)], 0.95)
    torch.cuda.synchronize()
    scale = ref.abs().max().item()
    torch.testing.assert_close(got, ref, rtol=1e-4, atol=1e-5 * scale)
    assert torch.equal(got, got.t())
