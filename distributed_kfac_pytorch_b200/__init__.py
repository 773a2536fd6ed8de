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

__version__ = '0.1.0'
__all__ = ['KFAC', 'CommMethod', 'KFACParamScheduler', 'modules', 'utils', 'comm', 'layers']
