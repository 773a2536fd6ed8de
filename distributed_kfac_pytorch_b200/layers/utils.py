"""Stand-alone forms of the per-layer math helpers of the reference's ``kfac/layers/utils.py``
(same names, arguments and error behaviour; file:line cites are into /root/reference).

``KFACLayer`` itself never calls these: its factor / inverse / precondition stages are batched
over all layers in ``libkfac_b200.so``.  They exist for user code that imports
``kfac.layers.utils`` directly, and every numerical one is a call into the same library (CUDA
tensors only, fp32 results); the list / shape helpers are plain tensor bookkeeping.
"""
import torch

from .._native import KFB_SEG_ROWS, FactorJob, Segment, get_ops


def append_bias_ones(tensor):
    """Append a column of ones on the last dimension (kfac/layers/utils.py:4-11)."""
    shape = list(tensor.shape[:-1]) + [1]
    return torch.cat([tensor, tensor.new_ones(shape)], dim=-1)


def _padded(rows, cols, like):
    pitch = (cols + 3) // 4 * 4
    return torch.empty((rows, pitch), device=like.device, dtype=torch.float32)[:, :cols]


def get_cov(a, b=None, scale=None):
    """Second moment ``a^T a / scale`` (symmetrised), or ``a^T b / scale``
    (kfac/layers/utils.py:13-43).  ``scale`` defaults to the row count."""
    if len(a.shape) != 2:
        raise ValueError('Input tensor must have 2 dimensions.')
    if b is not None and a.shape != b.shape:
        raise ValueError('Input tensors must have same shape. Got tensors of '
                         'shape {} and {}.'.format(a.shape, b.shape))
    if scale is None:
        scale = a.size(0)
    ops = get_ops()
    if b is None:
        # the factor SYRK with decay 0: state <- X^T X * weight, both triangles
        a = a if a.stride(1) == 1 else a.contiguous()
        out = _padded(a.shape[1], a.shape[1], a)
        seg = Segment(KFB_SEG_ROWS, a, n=a.shape[0], c=a.shape[1], weight=1.0 / float(scale), ld=a.stride(0))
        ops.factor_update([FactorJob(a.shape[1], [seg], out, True)], 0.0)
        return out
    at = _padded(a.shape[1], a.shape[0], a)
    bt = _padded(b.shape[1], b.shape[0], b)
    ops.transpose(a.float(), at)
    ops.transpose((b / scale).float(), bt)
    out = _padded(a.shape[1], b.shape[1], a)
    ops.gemm_tn(at, bt, out)
    return out


def get_eigendecomp(tensor, clip=0.0, concat=True, symmetric=True):
    """Eigendecomposition of a symmetric block by cuSOLVER syevd (kfac/layers/utils.py:45-74).
    Eigenvalues below ``clip`` are raised to it (the library clips at 0 itself)."""
    if not symmetric:
        raise NotImplementedError('only symmetric blocks are supported')
    d = tensor.shape[0]
    factor = _padded(d, d, tensor)
    factor.copy_(tensor)
    E, Q = _padded(d, d, tensor), _padded(d, d, tensor)
    w = torch.empty(d, device=tensor.device, dtype=torch.float32)
    info = torch.zeros(1, device=tensor.device, dtype=torch.int32)
    if clip is None:
        raise NotImplementedError('eigenvalues are always clipped at 0 by the native solver')
    get_ops().syevd(factor, E, Q, w, info)
    if clip > 0.0:
        w = torch.clamp_min(w, clip)
    if concat:
        return torch.cat([Q, w.unsqueeze(-1)], -1)
    return Q, w


def get_inverse(tensor, damping=None, symmetric=True):
    """``(tensor + damping I)^-1`` by Cholesky (kfac/layers/utils.py:76-96)."""
    if not symmetric:
        raise NotImplementedError('only symmetric blocks are supported')
    d = tensor.shape[0]
    factor = _padded(d, d, tensor)
    factor.copy_(tensor)
    out = _padded(d, d, tensor)
    info = torch.zeros(1, device=tensor.device, dtype=torch.int32)
    get_ops().cholesky_inverse(factor, 0.0 if damping is None else damping, out, info)
    return out


def get_elementwise_inverse(vector, damping=None):
    """Reciprocal of every non-zero element (kfac/layers/utils.py:98-105)."""
    if damping is not None:
        vector = vector + damping
    return torch.where(vector != 0.0, torch.reciprocal(vector), vector)


def reshape_data(data_list, batch_first=True, collapse_dims=False):
    """Concatenate captures along the batch dimension (kfac/layers/utils.py:107-124)."""
    d = torch.cat(data_list, dim=int(not batch_first))
    if collapse_dims and len(d.shape) > 2:
        d = d.view(-1, d.shape[-1])
    return d


def get_triu(tensor):
    """Flattened upper triangle, row by row (kfac/layers/utils.py:126-135)."""
    if len(tensor.shape) != 2:
        raise ValueError('triu(tensor) requires tensor to be 2 dimensional')
    if tensor.shape[0] > tensor.shape[1]:
        raise ValueError('tensor cannot have more rows than columns')
    if tensor.shape[0] != tensor.shape[1]:
        raise NotImplementedError('the packed-triangle kernels take square tensors')
    d = tensor.shape[0]
    src = tensor if tensor.stride(1) == 1 and tensor.dtype == torch.float32 else tensor.float().contiguous()
    tri = torch.empty(d * (d + 1) // 2, device=tensor.device, dtype=torch.float32)
    get_ops().tri_pack([(tri, src)])
    return tri


def fill_triu(shape, triu_tensor):
    """Symmetric tensor rebuilt from its flattened upper triangle (kfac/layers/utils.py:137-162)."""
    if len(shape) != 2:
        raise ValueError('shape must be 2 dimensional')
    rows, cols = shape
    if rows != cols:
        raise NotImplementedError('the packed-triangle kernels take square tensors')
    out = _padded(rows, cols, triu_tensor)
    get_ops().tri_unpack([(triu_tensor.contiguous(), out)], 1.0)
    return out


def update_running_avg(new, current, alpha=1.0):
    """In-place ``current = alpha*current + (1-alpha)*new`` in the reference's operation order
    (kfac/layers/utils.py:164-178); a no-op for ``alpha == 1``."""
    if alpha != 1:
        current *= alpha / (1 - alpha)
        current += new
        current *= (1 - alpha)
