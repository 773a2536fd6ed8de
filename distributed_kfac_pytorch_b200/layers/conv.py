"""Conv2d K-FAC layer: describes the captured NCHW tensors; the kernels do the
im2col implicitly (reference: kfac/layers/conv.py in /root/reference)."""
from .._native import KFB_SEG_IM2COL, KFB_SEG_NCHW, Segment
from .base import KFACLayer


class Conv2dLayer(KFACLayer):
    def __init__(self, *args, **kwargs):
        super(Conv2dLayer, self).__init__(*args, **kwargs)
        self.has_bias = self.module.bias is not None
        if not self.batch_first:
            raise ValueError('Conv2D layer must use batch_first=True')
        m = self.module
        if m.groups != 1 or tuple(m.dilation) != (1, 1) or isinstance(m.padding, str):
            raise NotImplementedError('K-FAC B200 supports groups=1, dilation=1, numeric padding')

    def _out_hw(self, h, w):
        m = self.module
        oh = (h + 2 * m.padding[0] - m.kernel_size[0]) // m.stride[0] + 1
        ow = (w + 2 * m.padding[1] - m.kernel_size[1]) // m.stride[1] + 1
        return oh, ow

    def _a_segments(self, a_inputs):
        """A = U^T U / (N S^3): rows pre-divided by S = oh*ow, then by the row count
        (kfac/layers/conv.py:24-34 + kfac/layers/utils.py:36-41)."""
        m = self.module
        kh, kw = m.kernel_size
        rows = 0
        shapes = []
        for a in a_inputs:
            n, c, h, w = a.shape
            oh, ow = self._out_hw(h, w)
            shapes.append((n, c, h, w, oh * ow))
            rows += n * oh * ow
        segs = []
        for a, (n, c, h, w, s) in zip(a_inputs, shapes):
            weight = 1.0 / (float(s) * float(s) * float(rows))
            pointwise = (kh == 1 and kw == 1 and tuple(m.stride) == (1, 1)
                         and tuple(m.padding) == (0, 0) and not self.has_bias)
            kind = KFB_SEG_NCHW if pointwise else KFB_SEG_IM2COL
            segs.append(Segment(kind, a.contiguous(), n, c, weight, has_bias=self.has_bias,
                                h=h, w=w, kernel=(kh, kw), stride=tuple(m.stride),
                                padding=tuple(m.padding)))
        c = a_inputs[0].shape[1]
        return c * kh * kw + int(self.has_bias), segs

    def _g_segments(self, g_outputs):
        """G = g^T g / (N S^3) with g read straight from NCHW (kfac/layers/conv.py:36-48)."""
        rows = sum(g.shape[0] * g.shape[2] * g.shape[3] for g in g_outputs)
        segs = []
        for g in g_outputs:
            n, c, h, w = g.shape
            s = h * w
            segs.append(Segment(KFB_SEG_NCHW, g.contiguous(), n, c,
                                1.0 / (float(s) * float(s) * float(rows)), h=h, w=w))
        return g_outputs[0].shape[1], segs
