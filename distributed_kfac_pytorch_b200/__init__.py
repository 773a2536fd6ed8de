"""B200-native K-FAC preconditioner with the API of MLHPC/Distributed_KFAC_Pytorch.

    from distributed_kfac_pytorch_b200 import KFAC, CommMethod, KFACParamScheduler

(or ``import kfac`` through the alias package at the repository root).  The Python
host code calls hand-written sm_100a CUDA kernels in ``libkfac_b200.so`` through the
C ABI declared in ``include/kfac_b200.h``; there is no CPU fallback."""
from .preconditioner import CommMethod, KFAC
from .scheduler import KFACParamScheduler
from . import modules
from . import utils
from . import comm
from . import layers



def set_mma_mode(mode):
    """Precision of the GEMM-shaped kernels, process-wide (no counterpart in the reference, whose math is
    fp32 throughout): 1 = single-pass TF32 tensor cores (default; inside the parity tolerance), 2 = the same
    with error-compensated 3xTF32 products in the preconditioning chain (gradients as accurate as fp32, about
    1 ms more per step on ResNet-50), 0 = exact fp32 FMA, no tensor cores.  Also: environment KFB_MMA_MODE."""
    from ._native import get_ops
    get_ops().set_mma_mode(mode)


def get_mma_mode():
    from ._native import get_ops
    return get_ops().get_mma_mode()


__version__ = '0.1.0'
__all__ = ['KFAC', 'CommMethod', 'KFACParamScheduler', 'modules', 'utils', 'comm', 'layers', 'set_mma_mode',
           'get_mma_mode']
