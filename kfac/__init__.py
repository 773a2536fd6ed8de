"""Drop-in alias: ``import kfac`` resolves to the B200-native implementation, so the
reference's example scripts (``kfac.KFAC``, ``kfac.CommMethod``, ``kfac.comm.backend``,
``kfac.utils``, ``kfac.modules.LSTM`` ...) run unmodified."""
import sys as _sys

import distributed_kfac_pytorch_b200 as _impl
from distributed_kfac_pytorch_b200 import (KFAC, CommMethod, KFACParamScheduler, comm, get_mma_mode, layers,
                                           modules, set_mma_mode, utils)

for _name in ('comm', 'layers', 'modules', 'utils', 'preconditioner', 'scheduler'):
    _sys.modules[__name__ + '.' + _name] = getattr(_impl, _name)
_sys.modules[__name__ + '.modules.lstm'] = _impl.modules.lstm
_sys.modules[__name__ + '.layers.utils'] = _impl.layers.utils
for _sub in ('base', 'conv', 'linear', 'embedding'):
    _sys.modules[__name__ + '.layers.' + _sub] = getattr(_impl.layers, _sub)

__all__ = ['KFAC', 'CommMethod', 'KFACParamScheduler', 'comm', 'layers', 'modules', 'utils', 'set_mma_mode',
           'get_mma_mode']
