"""Which combinations of split outputs does cuDevResourceGenerateDesc accept?"""
import torch
from cuda.bindings import driver as drv


def chk(res):
    if int(res[0]) != 0:
        raise RuntimeError('driver call failed: %s' % (res[0],))
    return res[1] if len(res) == 2 else res[1:]


torch.zeros(1, device='cuda:0')
dev = chk(drv.cuDeviceGet(0))
sm = chk(drv.cuDeviceGetDevResource(dev, drv.CUdevResourceType.CU_DEV_RESOURCE_TYPE_SM))
for ask in (32, 18, 15):
    try:
        groups, nb, rem = chk(drv.cuDevSmResourceSplitByCount(ask, sm, 0, 8))
        print('ask %d groups of 8: got %d, sizes %s, remainder %d' % (ask, nb, [g.sm.smCount for g in groups[:nb]], rem.sm.smCount), flush=True)
    except Exception as e:
        print('ask %d failed: %r' % (ask, e), flush=True)
groups, nb, rem = chk(drv.cuDevSmResourceSplitByCount(18, sm, 0, 8))
groups = list(groups[:nb])


def try_desc(label, res):
    try:
        desc = chk(drv.cuDevResourceGenerateDesc(res, len(res)))
        gctx = chk(drv.cuGreenCtxCreate(desc, dev, int(drv.CUgreenCtxCreate_flags.CU_GREEN_CTX_DEFAULT_STREAM)))
        st = chk(drv.cuGreenCtxStreamCreate(gctx, int(drv.CUstream_flags.CU_STREAM_NON_BLOCKING), 0))
        got = chk(drv.cuGreenCtxGetDevResource(gctx, drv.CUdevResourceType.CU_DEV_RESOURCE_TYPE_SM))
        print('%-40s ok: context has %d SMs' % (label, got.sm.smCount), flush=True)
        return gctx, st
    except Exception as e:
        print('%-40s FAILED: %r' % (label, e), flush=True)


keep = []
keep.append(try_desc('one group', [groups[0]]))
keep.append(try_desc('two groups', [groups[1], groups[2]]))
keep.append(try_desc('six groups', groups[3:9]))
keep.append(try_desc('remainder alone', [rem]))
keep.append(try_desc('group + remainder', [groups[9], rem]))
keep.append(try_desc('five groups + remainder', groups[10:15] + [rem]))
# other split granularities
for per in (16, 24, 32, 48, 64):
    try:
        g2, n2, r2 = chk(drv.cuDevSmResourceSplitByCount(18, sm, 0, per))
        print('split by %d: %d groups %s + %d' % (per, n2, [g.sm.smCount for g in g2[:n2]], r2.sm.smCount), flush=True)
    except Exception as e:
        print('split by %d failed: %r' % (per, e), flush=True)
