"""CPU restatement of the reference K-FAC hot path.  TEST INFRASTRUCTURE ONLY.

This module restates, in plain torch-on-CPU fp32 arithmetic, the algorithm of
MLHPC/Distributed_KFAC_Pytorch's ``KFAC.step()`` (reference
``kfac/preconditioner.py:472-523``) and everything it calls.  It exists to CHECK
the CUDA product path; nothing under ``distributed_kfac_pytorch_b200/`` may import
it.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs use it.

Why torch and not numpy: every FLOP of the reference is a torch CPU op
(``setup.py:19-21`` pins only ``torch``), so restating in the same numeric
library lets the oracle be pinned to the real reference at ~1e-6.

Parity status: PINNED.  ``tests/golden/make_golden.py`` runs the unmodified
reference (``/root/reference``, with the ``torch.symeig`` shim of SURVEY §8c) on
seeded inputs and stores factors / gradients / worker assignments under
``tests/golden/``; ``tests/test_oracle_vs_golden.py`` checks every function here
against those fixtures and against the reference's own known-answer tests
(``tests/load_balance.py``, ``tests/worker_allocator.py``).

All ``file:line`` citations are into ``/root/reference``.
"""
import math

import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------
# factor math                                   (kfac/layers/utils.py, conv.py, linear.py)
# --------------------------------------------------------------------------


def with_ones_column(t):
    """Append a column of ones on the last dim (kfac/layers/utils.py:4-11)."""
    return torch.cat([t, t.new_ones(*t.shape[:-1], 1)], dim=-1)


def second_moment(a, scale=None):
    """``a^T (a/scale)`` then symmetrise (kfac/layers/utils.py:13-43, b=None branch).

    scale defaults to the row count."""
    if a.dim() != 2:
        raise ValueError('Input tensor must have 2 dimensions.')
    if scale is None:
        scale = a.size(0)
    c = a.t() @ (a / scale)
    return (c + c.t()) / 2.0


def running_average_(new, cur, alpha):
    """In-place EMA in the reference's operation order (kfac/layers/utils.py:164-178).

    ``alpha == 1`` leaves ``cur`` untouched (App. A quirk)."""
    if alpha != 1:
        cur *= alpha / (1 - alpha)
        cur += new
        cur *= (1 - alpha)
    return cur


def unfold_patches(x, kernel_size, stride, padding):
    """im2col: (N,C,H,W) -> (N, oh, ow, C*kh*kw), column order (c, kh, kw).

    Restates kfac/layers/conv.py:50-70 with ``F.unfold`` (same ordering)."""
    n = x.shape[0]
    kh, kw = kernel_size
    oh = (x.shape[2] + 2 * padding[0] - kh) // stride[0] + 1
    ow = (x.shape[3] + 2 * padding[1] - kw) // stride[1] + 1
    cols = F.unfold(x, kernel_size, padding=padding, stride=stride)  # (N, C*kh*kw, oh*ow)
    return cols.transpose(1, 2).reshape(n, oh, ow, -1)


def conv2d_A(a_inputs, kernel_size, stride, padding, has_bias):
    """Conv2d input-covariance factor (kfac/layers/conv.py:24-34).

    Rows are divided by S=oh*ow, then get_cov divides by the row count N*S:
    A = U^T U / (N S^3) with the ones column (if bias) appended BEFORE the /S."""
    rows = []
    for a in a_inputs:
        p = unfold_patches(a, kernel_size, stride, padding)
        spatial = p.size(1) * p.size(2)
        p = p.reshape(-1, p.size(-1))
        if has_bias:
            p = with_ones_column(p)
        rows.append(p / spatial)
    return second_moment(torch.cat(rows, dim=0))


def conv2d_G(g_outputs):
    """Conv2d output-gradient factor (kfac/layers/conv.py:36-48)."""
    rows = []
    for g in g_outputs:
        spatial = g.size(2) * g.size(3)
        g2 = g.permute(0, 2, 3, 1).reshape(-1, g.size(1))
        rows.append(g2 / spatial)
    return second_moment(torch.cat(rows, dim=0))


def _collapse(data_list, batch_first):
    """kfac/layers/utils.py:107-124 with collapse_dims=True."""
    d = torch.cat(data_list, dim=int(not batch_first))
    if d.dim() > 2:
        d = d.reshape(-1, d.shape[-1])
    return d


def linear_A(a_inputs, has_bias, batch_first=True):
    """kfac/layers/linear.py:12-18."""
    a = _collapse(a_inputs, batch_first)
    if has_bias:
        a = with_ones_column(a)
    return second_moment(a)


def linear_G(g_outputs, batch_first=True):
    """kfac/layers/linear.py:20-24."""
    return second_moment(_collapse(g_outputs, batch_first))


def linear_multi_A(a_inputs, has_bias, batch_first=True):
    """Sum over calls of the per-call covariance (kfac/layers/linear.py:37-47)."""
    total = None
    for a in a_inputs:
        c = linear_A([a], has_bias, batch_first)
        total = c if total is None else total + c
    return total


def linear_multi_G(g_outputs, batch_first=True):
    """kfac/layers/linear.py:49-59."""
    total = None
    for g in g_outputs:
        c = linear_G([g], batch_first)
        total = c if total is None else total + c
    return total


def factor_ema(new, cur, alpha):
    """Running-average update incl. the identity seed (kfac/layers/base.py:381-417)."""
    if cur is None:
        cur = torch.eye(new.shape[0], dtype=new.dtype)
    return running_average_(new, cur, alpha)


# --------------------------------------------------------------------------
# inverses                                             (kfac/layers/utils.py:45-96, base.py:198-308)
# --------------------------------------------------------------------------


def eigh_clipped(factor):
    """``Q, d`` of a symmetric factor, eigenvalues clipped at 0.

    kfac/layers/utils.py:45-74 calls the removed ``torch.symeig``; the harness shim
     and this oracle use ``torch.linalg.eigh`` (LAPACK syevd)."""
    d, q = torch.linalg.eigh(factor.to(torch.float32))
    d = torch.clamp(d, min=0.0)
    return q.contiguous(), d


def outer_recip(dG, dA, damping):
    """``1 / (dG dA^T + damping)`` (kfac/layers/base.py:305-306)."""
    return 1.0 / (dG.unsqueeze(1) * dA.unsqueeze(0) + damping)


def damped_inverse(factor, damping):
    """Cholesky inverse of ``factor + damping I`` (kfac/layers/utils.py:76-96)."""
    f = factor.to(torch.float32) + damping * torch.eye(factor.shape[0])
    return torch.cholesky_inverse(torch.linalg.cholesky(f))


# --------------------------------------------------------------------------
# preconditioning                                     (kfac/layers/base.py:310-362, 459-475)
# --------------------------------------------------------------------------


def grad_matrix(weight_grad, bias_grad=None):
    """``[dW.view(out,-1) | db]`` (kfac/layers/base.py:310-319, conv.py:17-22)."""
    g = weight_grad.reshape(weight_grad.size(0), -1)
    if bias_grad is not None:
        g = torch.cat([g, bias_grad.reshape(-1, 1)], dim=1)
    return g


def precondition_eigen(g, QA, QG, dGdA=None, dA=None, dG=None, damping=0.001):
    """kfac/layers/base.py:459-470."""
    v1 = QG.t() @ g @ QA
    if dGdA is not None:
        v2 = v1 * dGdA
    else:
        v2 = v1 / (dG.unsqueeze(1) * dA.unsqueeze(0) + damping)
    return QG @ v2 @ QA.t()


def precondition_inv(g, A_inv, G_inv):
    """kfac/layers/base.py:472-475."""
    return G_inv @ g @ A_inv


def split_grad(v, weight_shape, has_bias):
    """Split the preconditioned matrix back (kfac/layers/base.py:355-362)."""
    if has_bias:
        return [v[:, :-1].reshape(weight_shape), v[:, -1:].reshape(-1)]
    return [v.reshape(weight_shape)]


def kl_clip_scale(precond, grads, lr, kl_clip):
    """``min(1, sqrt(kl_clip/|sum v*g*lr^2|))`` (kfac/preconditioner.py:661-682).

    ``precond`` / ``grads`` are flat lists of matching tensors (weight, bias, ...).
    Returns None when the sum is exactly 0, like the reference."""
    vg = 0.0
    for v, g in zip(precond, grads):
        vg += (v * g * lr ** 2).sum().item()
    if vg == 0.0:
        return None
    return min(1.0, math.sqrt(kl_clip / abs(vg)))


# --------------------------------------------------------------------------
# work placement  (integer work: bit-exact)      (kfac/utils.py:59-196, preconditioner.py:616-659)
# --------------------------------------------------------------------------


def load_balance(n_workers, work):
    """Greedy longest-processing-time placement (kfac/utils.py:169-196).

    Stable sort by descending cost, each item to the least-loaded worker with
    the lowest index winning ties."""
    if not n_workers > 0:
        raise ValueError('n_workers must be > 0')
    if len(work) == 0:
        raise ValueError('work cannot be an empty list')
    order = sorted(range(len(work)), key=lambda i: -work[i])  # stable
    load = [0] * n_workers
    out = [0] * len(work)
    for i in order:
        w = min(range(n_workers), key=lambda r: (load[r], r))
        load[w] += work[i]
        out[i] = w
    return out


def partition_grad_ranks(size, grad_workers):
    """kfac/utils.py:150-153."""
    return [list(range(i, size, grad_workers)) for i in range(grad_workers)]


def partition_inv_ranks(size, grad_workers):
    """kfac/utils.py:156-159."""
    return [list(range(i, min(i + grad_workers, size))) for i in range(0, size, grad_workers)]


def assign_workers(a_dims, g_dims, world, grad_worker_fraction, strategy='compute',
                   distribute_layer_factors=False):
    """Restates KFAC._assign_workers + WorkerAllocator (preconditioner.py:616-659,
    kfac/utils.py:59-147).  Returns one dict per layer:
    ``{'a_rank','g_rank','inv_ranks' (= gradient workers), 'grad_src_group': [(src, ranks) per rank]}``."""
    p = 3 if strategy == 'compute' else 2
    a_cost = [n ** p for n in a_dims]
    g_cost = [n ** p for n in g_dims]
    if distribute_layer_factors:
        locs = load_balance(world, a_cost + g_cost)
        a_locs, g_locs = locs[:len(a_cost)], locs[len(a_cost):]
    else:
        locs = load_balance(world, [x + y for x, y in zip(a_cost, g_cost)])
        a_locs, g_locs = locs, locs
    grad_workers = max(1, round(world * grad_worker_fraction))
    if world % grad_workers != 0:
        raise ValueError('compute_grad_fraction must produce equally size groups')
    grad_parts = partition_grad_ranks(world, grad_workers)
    inv_parts = partition_inv_ranks(world, grad_workers)

    def part_of(rank, parts):
        return next(p_ for p_ in parts if rank in p_)

    layers = []
    for a_rank, g_rank in zip(a_locs, g_locs):
        inv_ranks = part_of(a_rank, inv_parts)
        pairs = []
        for r in range(world):
            grp = part_of(r, grad_parts)
            src = next(x for x in grp if x in inv_ranks)
            pairs.append((src, grp))
        layers.append({'a_rank': a_rank, 'g_rank': g_rank, 'inv_ranks': inv_ranks,
                       'grad_src_group': pairs})
    return layers


# --------------------------------------------------------------------------
# whole step, single process (optionally emulating W data-parallel ranks)
# --------------------------------------------------------------------------


class LayerSpec(object):
    """What the oracle needs to know about one registered module.

    kind: 'conv2d' | 'linear' | 'linear_multi'."""

    def __init__(self, kind, weight_shape, has_bias, kernel_size=None, stride=None,
                 padding=None, batch_first=True):
        self.kind = kind
        self.weight_shape = tuple(weight_shape)
        self.has_bias = has_bias
        self.kernel_size = kernel_size
        self.stride = stride
        self.padding = padding
        self.batch_first = batch_first

    def A(self, a_inputs):
        if self.kind == 'conv2d':
            return conv2d_A(a_inputs, self.kernel_size, self.stride, self.padding, self.has_bias)
        if self.kind == 'linear':
            return linear_A(a_inputs, self.has_bias, self.batch_first)
        return linear_multi_A(a_inputs, self.has_bias, self.batch_first)

    def G(self, g_outputs):
        if self.kind == 'conv2d':
            return conv2d_G(g_outputs)
        if self.kind == 'linear':
            return linear_G(g_outputs, self.batch_first)
        return linear_multi_G(g_outputs, self.batch_first)


class OracleKFAC(object):
    """Single-process restatement of ``KFAC.step()`` (kfac/preconditioner.py:472-523).

    ``step(captures, grads)``:
      captures: list over emulated ranks of list over layers of ``(a_inputs, g_outputs)``
                (only read on factor-update steps; may be None otherwise)
      grads:    list over layers of ``(weight_grad, bias_grad|None)``, already averaged
                over ranks (the reference requires that, preconditioner.py:479-481)
    returns the list over layers of ``[weight_grad', bias_grad']`` after K-FAC.
    Emulated ranks reproduce allreduce_factors (each rank's EMA, then the mean:
    base.py:114-127 + comm.py:254-262)."""

    def __init__(self, specs, damping=0.001, factor_decay=0.95, factor_update_freq=10,
                 inv_update_freq=100, kl_clip=0.001, lr=0.1, use_eigen_decomp=True):
        self.specs = specs
        self.damping = damping
        self.factor_decay = factor_decay
        self.factor_update_freq = factor_update_freq
        self.inv_update_freq = inv_update_freq
        self.kl_clip = kl_clip
        self.lr = lr
        self.use_eigen_decomp = use_eigen_decomp
        self.steps = 0
        self.A = [None] * len(specs)
        self.G = [None] * len(specs)
        self.inv = [None] * len(specs)

    def update_factors(self, captures):
        world = len(captures)
        for i, spec in enumerate(self.specs):
            accA, accG = None, None
            for r in range(world):
                a_in, g_out = captures[r][i]
                curA = None if self.A[i] is None else self.A[i].clone()
                curG = None if self.G[i] is None else self.G[i].clone()
                newA = factor_ema(spec.A(a_in), curA, self.factor_decay)
                newG = factor_ema(spec.G(g_out), curG, self.factor_decay)
                accA = newA if accA is None else accA + newA
                accG = newG if accG is None else accG + newG
            self.A[i] = accA / world if world > 1 else accA
            self.G[i] = accG / world if world > 1 else accG

    def update_inverses(self):
        for i in range(len(self.specs)):
            if self.use_eigen_decomp:
                QA, dA = eigh_clipped(self.A[i])
                QG, dG = eigh_clipped(self.G[i])
                self.inv[i] = (QA, QG, outer_recip(dG, dA, self.damping))
            else:
                self.inv[i] = (damped_inverse(self.A[i], self.damping),
                               damped_inverse(self.G[i], self.damping))

    def precondition(self, grads):
        out = []
        for i, spec in enumerate(self.specs):
            w, b = grads[i]
            g = grad_matrix(w, b if spec.has_bias else None)
            if self.use_eigen_decomp:
                QA, QG, dGdA = self.inv[i]
                v = precondition_eigen(g, QA, QG, dGdA=dGdA)
            else:
                v = precondition_inv(g, *self.inv[i])
            out.append(split_grad(v, spec.weight_shape, spec.has_bias))
        return out

    def step(self, captures, grads):
        if self.steps % self.factor_update_freq == 0:
            self.update_factors(captures)
        if self.steps % self.inv_update_freq == 0:
            self.update_inverses()
        v = self.precondition(grads)
        scale = None
        if self.kl_clip is not None:
            flat_v, flat_g = [], []
            for i, spec in enumerate(self.specs):
                flat_v.append(v[i][0])
                flat_g.append(grads[i][0])
                if spec.has_bias:
                    flat_v.append(v[i][1])
                    flat_g.append(grads[i][1])
            scale = kl_clip_scale(flat_v, flat_g, self.lr, self.kl_clip)
        if scale is not None:
            v = [[scale * t for t in layer] for layer in v]
        self.steps += 1
        return v
