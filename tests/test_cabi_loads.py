"""CPU-side checks of the C-ABI boundary: the shared library loads without a GPU and
exports every symbol declared in include/kfac_b200.h; argument errors are reported through
kfb_last_error(); the product path refuses to run without CUDA tensors."""
import os
import re

import pytest
import torch

from conftest import ROOT
from distributed_kfac_pytorch_b200 import _native


def _declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'kfac_b200.h')).read()
    return sorted(set(re.findall(r'KFB_API [\w\* ]+?\b(kfb_\w+)\(', text)))


@pytest.fixture(scope='module')
def lib():
    if not os.path.exists(_native.library_path()):
        import __graft_entry__
        __graft_entry__.build()
    return _native.load_library()


def test_header_and_binding_agree(lib):
    declared = _declared_symbols()
    assert len(declared) >= 16
    assert declared == sorted(_native.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.kfb_version() == 2


def test_errors_without_context(lib):
    if torch.cuda.is_available():
        pytest.skip('context exists on a GPU box')
    rc = lib.kfb_tri_pack(None, 0, None)
    assert rc != 0
    assert b'kfb_init' in lib.kfb_last_error()
    assert lib.kfb_set_mma_mode(7) != 0
    assert b'mma mode' in lib.kfb_last_error()
    prev = lib.kfb_get_mma_mode()
    for mode in (0, 2, 1):               # SIMT fp32, TF32 + 3xTF32 chain, TF32
        assert lib.kfb_set_mma_mode(mode) == 0 and lib.kfb_get_mma_mode() == mode
    assert lib.kfb_set_mma_mode(prev) == 0
    # the workspace promise is bookkeeping only: no device needed, bad pointers refused
    assert lib.kfb_factor_workspace_dedicate(None, 0) != 0
    assert lib.kfb_factor_workspace_dedicate(4096, 1 << 20) == 0 and lib.kfb_factor_workspace_release(4096) == 0
    assert lib.kfb_factor_workspace_release(4096) == 0      # releasing twice is harmless
    assert lib.kfb_init(0) != 0          # no device here
    assert lib.kfb_last_error() != b''


def test_no_cpu_fallback(lib):
    _native.set_ops(None)
    ops = _native.get_ops()
    assert ops.name == 'cuda'
    a = torch.eye(4)
    with pytest.raises(_native.NativeError):
        ops.transpose(a, a.clone())


def test_syevd_lane_planner(lib):
    """Host logic of kfb_syevd_batch: lanes of different widths from the measured time table."""
    # ResNet-50 on one GPU, the factors that go to syevd (the rest is batched)
    singles = [4608] * 3 + [2304] * 6 + [2048] * 7 + [1152] * 4 + [576] * 3
    lane_of, sms, ms = _native.syevd_lane_plan(singles, 148, 6)
    _, one_sms, one_ms = _native.syevd_lane_plan(singles, 148, 1)
    assert one_sms == [148] and 650 < one_ms < 800
    assert 2 <= len(sms) <= 6 and sum(sms) <= 148 and sms == sorted(sms, reverse=True)
    assert all(w % 8 == 0 or i == 0 for i, w in enumerate(sms))     # units of 8 SMs; lane 0 takes the remainder too
    assert ms < 0.5 * one_ms
    assert len(lane_of) == len(singles) and set(lane_of) <= set(range(len(sms)))
    # the 4608-dim factors sit on the widest lanes, the 576s on the narrow ones
    assert max(lane_of[i] for i in range(3)) <= min(lane_of[i] for i in range(len(singles) - 3, len(singles)))
    # rank 0 of 8 holds one 4608 + one 512 layer: one full-width stream beats any partition
    lane_of, sms, ms = _native.syevd_lane_plan([4608, 512], 148, 6)
    assert sms == [148] and lane_of == [0, 0] and 90 < ms < 100
    # two lanes, mid-size mix: wide + narrow
    lane_of, sms, ms = _native.syevd_lane_plan([2304, 2048, 2048, 1152, 1024, 1024, 576], 148, 2)
    assert len(sms) == 2 and sms[0] > sms[1] and sum(sms) == 148
    assert lib.kfb_syevd_lane_plan(None, 0, 148, 2, None, None, None, None) != 0
