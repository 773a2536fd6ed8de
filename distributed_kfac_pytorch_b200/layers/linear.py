"""Linear K-FAC layers (reference: kfac/layers/linear.py in /root/reference)."""
from .._native import KFB_SEG_ROWS, Segment
from .base import KFACLayer


def _rows_segment(t, weight, has_bias):
    feat = t.shape[-1]
    t2 = t.reshape(-1, feat)
    if t2.stride(-1) != 1:
        t2 = t2.contiguous()
    ld = t2.stride(0) if t2.shape[0] > 1 else feat
    return Segment(KFB_SEG_ROWS, t2, t2.shape[0], feat, weight, has_bias=has_bias, ld=ld)


class LinearLayer(KFACLayer):
    def __init__(self, *args, **kwargs):
        super(LinearLayer, self).__init__(*args, **kwargs)
        self.has_bias = self.module.bias is not None

    def _segments(self, tensors, has_bias):
        """All captures are concatenated along the batch dim, every leading dim is
        collapsed, and the covariance is divided by the total row count
        (kfac/layers/linear.py:12-24, kfac/layers/utils.py:107-124)."""
        rows = sum(t.numel() // t.shape[-1] for t in tensors)
        return [_rows_segment(t, 1.0 / rows, has_bias) for t in tensors]

    def _a_segments(self, a_inputs):
        return a_inputs[0].shape[-1] + int(self.has_bias), self._segments(a_inputs, self.has_bias)

    def _g_segments(self, g_outputs):
        return g_outputs[0].shape[-1], self._segments(g_outputs, False)


class LinearMultiLayer(LinearLayer):
    """Linear called once per timestep: the factor is the SUM over calls of the per-call
    covariance, each divided by its own row count (kfac/layers/linear.py:27-59)."""

    _renormalise_dropped = False     # a dropped AMP capture just leaves the sum: every call has its own divisor

    def _segments(self, tensors, has_bias):
        return [_rows_segment(t, 1.0 / (t.numel() // t.shape[-1]), has_bias) for t in tensors]
