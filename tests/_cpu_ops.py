"""CPU stand-in for the native ops object, built on the oracle.  TESTS ONLY.

It lets the host logic (hooks, segment descriptions, placement, collectives over gloo,
state handling) be exercised in the CPU-only container.  It is installed with
``_native.set_ops(OracleOps())``; the product never imports it."""
import math

import torch

from oracle import kfac_oracle as O
from distributed_kfac_pytorch_b200 import _native as N


def segment_matrix(s):
    """Materialise X_s of a Segment (what the CUDA kernels read implicitly)."""
    t = s.tensor.float()
    if s.kind == N.KFB_SEG_ROWS:
        x = t.reshape(-1, s.c)
    elif s.kind == N.KFB_SEG_NCHW:
        x = t.reshape(s.n, s.c, s.h * s.w).permute(0, 2, 1).reshape(-1, s.c)
    else:
        x = O.unfold_patches(t, s.kernel, s.stride, s.padding)
        x = x.reshape(-1, x.shape[-1])
    if s.has_bias:
        x = O.with_ones_column(x)
    return x


def triu_pack(m):
    i = torch.triu_indices(m.shape[0], m.shape[0])
    return m[i[0], i[1]]


def triu_unpack(tri, d):
    out = tri.new_zeros(d, d)
    i = torch.triu_indices(d, d)
    out[i[0], i[1]] = tri
    out = out + out.t()
    out.diagonal().mul_(0.5)
    return out


class OracleOps(object):
    name = 'oracle-cpu'

    def __init__(self):
        self.calls = []

    def factor_update(self, jobs, alpha):
        self.calls.append(('factor_update', len(jobs)))
        self.capture_filter([j.filter for j in jobs if j.filter is not None])
        for j in jobs:
            f = None
            for s in j.segments:
                mul = 1.0 if s.dev_scale is None else float(s.dev_scale)
                if mul == 0.0:
                    continue
                x = segment_matrix(s)
                c = (s.weight * mul) * (x.t() @ x)
                f = c if f is None else f + c
            old = torch.eye(j.d) if j.seed_identity else j.state.clone()
            resc = 1.0 if j.dev_rescale is None else float(j.dev_rescale)
            if resc < 0 or f is None:
                new = old
            else:
                f = (f + f.t()) / 2 * resc
                new = O.running_average_(f, old, alpha)
            if j.write_state:
                j.state.copy_(new)
            if j.tri_out is not None:
                j.tri_out.copy_(triu_pack(new))

    def capture_filter(self, filters):
        for f in filters:
            kept = total = 0.0
            for k, (t, rows, scale) in enumerate(zip(f.tensors, f.rows, f.scales)):
                ok = bool(torch.isfinite(t).all())
                f.dev_scales[k] = 1.0 / float(scale) ** 2 if ok else 0.0
                total += rows
                kept += rows if ok else 0.0
            f.dev_rescale[0] = -1.0 if kept == 0 else (total / kept if f.renormalise else 1.0)

    def narrow(self, src, dst16):
        dst16.copy_(src.to(dst16.dtype))

    def widen(self, src16, dst):
        dst.copy_(src16.float())

    def tri_pack(self, pairs):
        self.calls.append(('tri_pack', len(pairs)))
        for tri, state in pairs:
            tri.copy_(triu_pack(state))

    def tri_unpack(self, pairs, scale=1.0):
        self.calls.append(('tri_unpack', len(pairs)))
        for tri, state in pairs:
            state.copy_(triu_unpack(tri, state.shape[0]) * scale)

    def syevd(self, factor, E, Q, w, info):
        q, d = O.eigh_clipped(factor.clone())
        Q.copy_(q)
        E.copy_(q.t())
        w.copy_(d)
        info.zero_()

    def syevd_batch(self, jobs):
        for job in jobs:
            self.syevd(*job)

    def cholesky_inverse(self, factor, damping, out, info):
        out.copy_(O.damped_inverse(factor.clone(), damping))
        info.zero_()

    def transpose(self, src, dst):
        dst.copy_(src.t())

    def outer_recip(self, dG, dA, damping, out):
        out.copy_(O.outer_recip(dG, dA, damping))

    def precondition(self, jobs, vg_acc):
        self.calls.append(('precondition', len(jobs)))
        for j in jobs:
            g = O.grad_matrix(j.weight_grad, j.bias_grad)
            j.gmat.copy_(g)
            if j.use_inverse:
                j.out.copy_(O.precondition_inv(g, j.EA, j.QG))
            else:
                v2 = (j.EG @ g @ j.EA.t()) * j.dGdA        # = QG^T g QA
                j.out.copy_(j.QG @ v2 @ j.QA.t())
            if vg_acc is not None:
                vg_acc += (j.out * g).sum()

    def dot_accumulate(self, items, vg_acc):
        for wg, bg, v in items:
            vg_acc += (v * O.grad_matrix(wg, bg)).sum()

    def scale_writeback(self, items, vg_acc, lr, kl_clip, fixed_scale=1.0):
        self.calls.append(('scale_writeback', len(items)))
        nu = fixed_scale
        if vg_acc is not None:
            vg = float(vg_acc.item()) * lr ** 2
            if vg != 0.0:
                nu = min(1.0, math.sqrt(kl_clip / abs(vg)))
        for wg, bg, v in items:
            n_w = v.shape[1] - (1 if bg is not None else 0)
            wg.copy_((nu * v[:, :n_w]).reshape(wg.shape))
            if bg is not None:
                bg.copy_(nu * v[:, n_w])
