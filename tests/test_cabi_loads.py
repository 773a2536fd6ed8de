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
    assert lib.kfb_version() == 1


def test_errors_without_context(lib):
    if torch.cuda.is_available():
        pytest.skip('context exists on a GPU box')
    rc = lib.kfb_tri_pack(None, 0, None)
    assert rc != 0
    assert b'kfb_init' in lib.kfb_last_error()
    assert lib.kfb_set_mma_mode(7) != 0
    assert b'mma mode' in lib.kfb_last_error()
    assert lib.kfb_init(0) != 0          # no device here
    assert lib.kfb_last_error() != b''


def test_no_cpu_fallback(lib):
    _native.set_ops(None)
    ops = _native.get_ops()
    assert ops.name == 'cuda'
    a = torch.eye(4)
    with pytest.raises(_native.NativeError):
        ops.transpose(a, a.clone())
