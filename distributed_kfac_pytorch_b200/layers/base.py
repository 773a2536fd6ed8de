"""Per-layer K-FAC state and math, B200 edition.

``KFACLayer`` keeps the reference's per-layer interface (``kfac/layers/base.py`` in
/root/reference: same method names, the ``state`` dict with keys
``A, G, QA, QG, dA, dG, dGdA``, the same error behaviour) but every tensor operation is
a call into libkfac_b200.so through ``_native.get_ops()``:

* the factors are never formed from an unfolded activation matrix: the layer only
  DESCRIBES its captured tensors (``_a_segments`` / ``_g_segments``) and the SYRK kernel
  reads them in place (implicit im2col), fusing the running average;
* eigenvectors are stored twice, ``QA`` (columns = eigenvectors, what the reference
  exposes) and its transpose ``EA``, so that the four precondition GEMMs are all
  "K-major x K-major" tensor-core GEMMs with no transposed operand;
* matrices live in row-pitch-padded buffers (pitch a multiple of 4 floats = 16 B, the
  TMA global-stride rule); ``state[...]`` are views of the logical shape.
"""
import warnings

import torch

from .. import comm
from .._native import KFB_SEG_ROWS, CaptureFilter, FactorJob, PrecondJob, get_ops


def _round4(n):
    return (n + 3) // 4 * 4


def alloc2d(rows, cols, device, dtype=torch.float32, zero=False):
    """(buffer [rows, pitch] contiguous, view [rows, cols])."""
    make = torch.zeros if zero else torch.empty
    buf = make((rows, _round4(cols)), device=device, dtype=dtype)
    return buf, buf[:, :cols]


class KFACLayer(object):
    def __init__(self,
                 module,
                 accumulate_data=True,
                 batch_first=True,
                 inv_dtype=None,
                 grad_scaler=None,
                 factor_dtype=None,
                 prediv_eigenvalues=True,
                 symmetry_aware_comm=False,
                 use_eigen_decomp=True):
        self.module = module
        self.accumulate_data = accumulate_data
        self.batch_first = batch_first
        self.grad_scaler = grad_scaler
        self.inv_dtype = inv_dtype
        self.factor_dtype = factor_dtype
        self.prediv_eigenvalues = prediv_eigenvalues
        self.symmetry_aware_comm = symmetry_aware_comm
        self.use_eigen_decomp = use_eigen_decomp
        self.eps = 1e-10
        # factor_dtype: captures are rounded to it before the factor math (kfac/layers/base.py:385,409);
        # inv_dtype: eigenvectors / eigenvalues / dGdA (or the explicit inverses) are rounded to it after the
        # fp32 solve and travel in it (base.py:435-441, 237-292).  Accumulation, the running averages
        # and the GEMM chain stay fp32 / TF32 here: a 16-bit type costs the reference precision in every
        # product, here only in what it stores.
        for name, dt in (('inv_dtype', inv_dtype), ('factor_dtype', factor_dtype)):
            if dt not in (None, torch.float32, torch.float16, torch.bfloat16):
                raise NotImplementedError('%s must be float32, float16 or bfloat16, got %s' % (name, dt))

        # Should be overridden by implementing class
        self.has_bias = False
        self.factors_are_symmetric = True

        self.a_inputs = []   # inputs accumulated from module hooks
        self.g_outputs = []  # outputs accumulated from module hooks
        self.state = {'A': None, 'G': None}
        self.preconditioned_gradient = None

        self.compute_A_inv_rank = None
        self.compute_G_inv_rank = None
        self.broadcast_A_inv_group = None
        self.broadcast_G_inv_group = None
        self.compute_grad_ranks = None
        self.broadcast_grad_groups = None
        self.keep_inv_copy = None

        # internal device buffers (padded) keyed like the state entries they back
        self._buf = {}
        self._E = {'A': None, 'G': None}      # transposed eigenvector matrices
        self._info = None                     # cuSOLVER devInfo for (A, G)
        self._job = None                      # persistent PrecondJob
        self._scratch = None                  # (gmat, V, T) views
        self._out = None                      # dg x da preconditioned gradient view
        self._tri_fresh = {'A': False, 'G': False}
        self._amp_scalars = []                # per-capture device scalars of the AMP capture filters since the last check

    _renormalise_dropped = True               # dropped AMP captures shrink the divisor (get_cov over what is left)

    def _inv16(self):
        return self.inv_dtype in (torch.float16, torch.bfloat16)

    def __repr__(self):
        return 'KFAC {}({})'.format(self.__class__.__name__, repr(self.module))

    # ------------------------------------------------------------------ checkpoint
    def state_dict(self, include_inverses=False):
        """Layout is contract (kfac/layers/base.py:47-60): full square fp32 A and G.  The entries
        are contiguous copies of the logical shape - the live tensors are views of row-pitch-padded
        buffers (or of a broadcast slab), which would drag their whole storage into a checkpoint
        and which the reference's load_state_dict + allreduce could not take."""
        keys = self.state.keys() if include_inverses else ('A', 'G')
        return {k: (self.state[k].clone(memory_format=torch.contiguous_format)
                    if torch.is_tensor(self.state[k]) else self.state[k]) for k in keys}

    def load_state_dict(self, state_dict):
        """kfac/layers/base.py:62-75, including the legacy A_factor/G_factor keys."""
        if 'A' not in state_dict or 'G' not in state_dict:
            if 'A_factor' in state_dict and 'G_factor' in state_dict:
                state_dict['A'] = state_dict.pop('A_factor')
                state_dict['G'] = state_dict.pop('G_factor')
            else:
                raise KeyError('KFACLayer state_dict must contain keys '
                               '"A" and "G"')
        device = self._device()
        self.state = {'A': None, 'G': None}
        self._buf.clear()
        self._job = None
        for key, value in state_dict.items():
            if value is None:
                self.state[key] = None
            elif value.dim() == 2:
                self._set2d(key, value.shape[0], value.shape[1], device).copy_(value)
            else:
                self.state[key] = value.to(device=device, dtype=torch.float32).clone()
        for key in ('A', 'G'):
            q = self.state.get('Q' + key)
            if q is not None:
                get_ops().transpose(q, self._alloc_E(key, q.shape[0], device))
        self._tri_fresh = {'A': False, 'G': False}

    # ------------------------------------------------------------------ assignment
    def assign_inverse_workers(self, compute_A_inv_rank, compute_G_inv_rank,
                               broadcast_A_inv_group, broadcast_G_inv_group):
        """kfac/layers/base.py:77-94."""
        if compute_A_inv_rank != compute_G_inv_rank and self.prediv_eigenvalues:
            raise ValueError('When precomputing 1 / (dG * dA.T + damping), '
                             'A and G inverse worker ranks must be equal. '
                             'I.e. distribute_layer_factors=False.')
        self.compute_A_inv_rank = compute_A_inv_rank
        self.compute_G_inv_rank = compute_G_inv_rank
        self.broadcast_A_inv_group = broadcast_A_inv_group
        self.broadcast_G_inv_group = broadcast_G_inv_group

    def assign_gradient_workers(self, compute_grad_ranks, broadcast_grad_groups):
        """kfac/layers/base.py:96-112."""
        if len(broadcast_grad_groups) != comm.backend.size():
            raise ValueError('len(broadcast_grad_groups) != world size')
        self.compute_grad_ranks = compute_grad_ranks
        self.broadcast_grad_groups = broadcast_grad_groups
        self.keep_inv_copy = comm.backend.rank() in self.compute_grad_ranks

    # ------------------------------------------------------------------ buffers
    def _device(self):
        return next(self.module.parameters()).device

    def _set2d(self, key, rows, cols, device, zero=False):
        buf, view = alloc2d(rows, cols, device, zero=zero)
        self._buf[key] = buf
        self.state[key] = view
        return view

    def _inject2d(self, key, buf, cols):
        """Back ``state[key]`` (or the preconditioned gradient for key 'out') with a buffer
        owned by somebody else: KFAC lays the buffers of all layers that share a broadcast
        source out in one slab so that the whole exchange is a single collective."""
        self._buf[key] = buf
        view = buf[:, :cols]
        if key == 'out':
            self._out = None
            self._set_out(view)
        else:
            self.state[key] = view
        return view

    def _alloc_E(self, key, d, device):
        buf, view = alloc2d(d, d, device)
        self._buf['E' + key] = buf
        self._E[key] = view
        return view

    # ------------------------------------------------------------------ capture
    def save_inputs(self, input):
        """Keep a REFERENCE to the activation (no copy), kfac/layers/base.py:364-369."""
        if self.accumulate_data:
            self.a_inputs.append(input[0].data)
        else:
            self.a_inputs = [input[0].data]

    def save_grad_outputs(self, grad_output):
        """kfac/layers/base.py:371-379."""
        g = grad_output[0].data
        if self.grad_scaler is not None:
            # torch's GradScaler keeps its scale in a device scalar; get_scale() would synchronise with
            # the host on every hook call.  The scale cannot change before K-FAC consumes the capture
            # (scaler.update() runs after the optimizer step), so a reference to the tensor is enough.
            scale = getattr(self.grad_scaler, '_scale', None)
            if not (torch.is_tensor(scale) and scale.is_cuda and scale.dtype == torch.float32
                    and scale.numel() == 1):
                scale = self.grad_scaler.get_scale()
            g = (g, scale)
        if self.accumulate_data:
            self.g_outputs.append(g)
        else:
            self.g_outputs = [g]

    # ------------------------------------------------------------------ factors
    def _a_segments(self, a_inputs):
        """Describe the captured inputs for the SYRK kernel: (d, [Segment])."""
        raise NotImplementedError

    def _g_segments(self, g_outputs):
        raise NotImplementedError

    def _as_f32(self, t):
        """Captures are handed to the SYRK as they are when it can read them (fp32, and the fp16 /
        bf16 tensors autocast produces: they are widened while being gathered, never copied).  A 16-bit
        ``factor_dtype`` rounds fp32 captures to it first, as the reference's ``x.to(factor_dtype)``."""
        if self.factor_dtype in (torch.float16, torch.bfloat16, torch.float32) and t.dtype != self.factor_dtype:
            return t.to(self.factor_dtype)
        if t.dtype not in (torch.float32, torch.float16, torch.bfloat16):
            t = t.float()
        return t

    def _factor_job(self, key, d, segments, alpha, tri_out):
        """One kfb_factor_job, allocating / seeding the running average on first use
        (identity seed: kfac/layers/base.py:388-389,415-416)."""
        seed = self.state[key] is None
        if seed:
            self._set2d(key, d, d, segments[0].tensor.device)
            if self.grad_scaler is not None and key == 'G':
                # the very first update may be dropped on the device (all captures inf / nan)
                self.state[key].zero_()
                self.state[key].diagonal().fill_(1.0)
        if alpha == 1:
            # update_running_avg is a no-op for alpha == 1 (kfac/layers/utils.py:175):
            # a fresh factor stays the identity
            if seed:
                s = self.state[key]
                s.zero_()
                s.diagonal().fill_(1.0)
            return None
        self._tri_fresh[key] = tri_out is not None
        return FactorJob(d, segments, self.state[key], seed, True, tri_out)

    def _A_job(self, alpha, tri_out=None):
        if len(self.a_inputs) == 0:
            return None
        inputs = [self._as_f32(x) for x in self.a_inputs]
        self.a_inputs = []                       # clear accumulated inputs (base.py:387)
        d, segments = self._a_segments(inputs)
        return self._factor_job('A', d, segments, alpha, tri_out)

    def _G_job(self, alpha, tri_out=None):
        outputs = self.g_outputs
        self.g_outputs = []                      # clear accumulated outputs (base.py:414)
        if len(outputs) == 0:
            return None
        if self.grad_scaler is None:
            d, segments = self._g_segments([self._as_f32(x) for x in outputs])
            return self._factor_job('G', d, segments, alpha, tri_out)
        # AMP: the captures carry the loss scale and may hold inf / nan.  The reference unscales them,
        # tests every capture on the host and drops the bad ones (kfac/layers/base.py:397-407: one
        # synchronisation per capture).  Here the kernel kfb_capture_filter does it on the device: it
        # writes 1/s^2 (G is quadratic in g: (g/s)^T (g/s) = g^T g / s^2) or 0 into one scalar per capture,
        # which multiplies that segment's weight inside the SYRK, and one scalar per factor that
        # renormalises by the rows that are left, or tells the kernels to leave the running average alone.
        d, segments = self._g_segments([self._as_f32(g) for g, _ in outputs])
        n = len(segments)
        scalars = torch.empty(n + 1, device=segments[0].tensor.device, dtype=torch.float32)
        rows = []
        for k, seg in enumerate(segments):
            seg.dev_scale = scalars[k:k + 1]
            rows.append(seg.n if seg.kind == KFB_SEG_ROWS else seg.n * seg.h * seg.w)
        job = self._factor_job('G', d, segments, alpha, tri_out)
        if job is not None:
            job.dev_rescale = scalars[n:]
            job.filter = CaptureFilter([seg.tensor for seg in segments], rows, [s for _, s in outputs],
                                       self._renormalise_dropped, scalars[:n], scalars[n:])
            self._amp_scalars.append(scalars[:n])
        return job

    def amp_captures_dropped(self):
        """Device bool tensor: did a G update since the last call drop a capture (None: no AMP update)."""
        if not self._amp_scalars:
            return None
        flag = torch.cat(self._amp_scalars).eq(0).any()
        self._amp_scalars = []
        return flag

    def update_A_factor(self, alpha=0.95):
        """Compute factor A and fold it into the running average (base.py:381-390)."""
        job = self._A_job(alpha)
        if job is not None:
            get_ops().factor_update([job], alpha)

    def update_G_factor(self, alpha=0.95):
        """Compute factor G and fold it into the running average (base.py:392-417)."""
        job = self._G_job(alpha)
        if job is not None:
            get_ops().factor_update([job], alpha)

    def allreduce_factors(self):
        """Per-layer allreduce of the full square factors (base.py:114-127).  KFAC's own
        ``allreduce_factors`` uses the packed-triangle fusion buffer instead; this entry
        point is kept for callers that drive layers by hand."""
        handles = []
        for key in ('A', 'G'):
            buf = self._buf[key]
            handles.append(comm.backend.allreduce(buf))
        return handles

    # ------------------------------------------------------------------ inverses
    def _check_inv_ready(self, key):
        if (self.compute_A_inv_rank if key == 'A' else self.compute_G_inv_rank) is None:
            raise ValueError('Workers have not been assigned to layer yet.')
        if self.keep_inv_copy is None:
            raise ValueError('Grad workers have not been assigned to layer yet.')
        if self.state[key] is None:
            raise RuntimeError('update_{0}_factor() must be called at least '
                               'once before calling compute_{0}_inv().'.format(key))

    def _ensure_inv_buffers(self, key):
        d = self.state[key].shape[0]
        dev = self.state[key].device
        if self._info is None:
            self._info = torch.zeros(2, device=dev, dtype=torch.int32)
        if not self.use_eigen_decomp:
            if key + '_inv' not in self.state:
                self._set2d(key + '_inv', d, d, dev)
            return
        if 'Q' + key not in self.state:
            self._set2d('Q' + key, d, d, dev)
            self._alloc_E(key, d, dev)
        if 'd' + key not in self.state:
            self.state['d' + key] = torch.empty(d, device=dev, dtype=torch.float32)
        if self.prediv_eigenvalues and 'dGdA' not in self.state \
                and self.state['A'] is not None and self.state['G'] is not None:
            self._set2d('dGdA', self.state['G'].shape[0], self.state['A'].shape[0], dev)
        if self._info is None:
            self._info = torch.zeros(2, device=dev, dtype=torch.int32)

    def _inv_job(self, key, ignore_rank=False):
        """Checks and buffers for one factor's eigendecomposition; returns the syevd job
        ``(factor, E, Q, w, info)`` if THIS rank computes it, else None."""
        self._check_inv_ready(key)
        mine = comm.backend.rank() == (self.compute_A_inv_rank if key == 'A'
                                       else self.compute_G_inv_rank)
        if self.keep_inv_copy or mine or ignore_rank:
            self._ensure_inv_buffers(key)
        if ignore_rank or mine:
            info = self._info[0:1] if key == 'A' else self._info[1:2]
            if not self.use_eigen_decomp:
                return (self.state[key], self.state[key + '_inv'], info)
            return (self.state[key], self._E[key], self.state['Q' + key], self.state['d' + key], info)
        return None

    def _run_inv_job(self, job, damping):
        if self.use_eigen_decomp:
            get_ops().syevd(*job)
        else:     # explicit damped inverse (kfac/layers/utils.py:76-96)
            factor, out, info = job
            get_ops().cholesky_inverse(factor, damping, out, info)

    def _round16(self, *keys):
        """Round the named buffers to inv_dtype in place (they stay fp32 tensors holding 16-bit values)."""
        ops = get_ops()
        for key in keys:
            t = self._buf.get(key)
            if t is None:
                t = self.state.get(key)
            if t is None:
                continue
            flat = t.reshape(-1) if t.is_contiguous() else None
            if flat is None:                       # a 1-D view of a slab is contiguous; 2-D buffers are whole
                raise RuntimeError('cannot round a non-contiguous buffer')
            tmp = torch.empty(flat.numel(), device=flat.device, dtype=self.inv_dtype)
            ops.narrow(flat, tmp)
            ops.widen(tmp, flat)

    def _finish_A_inv(self):
        if self._inv16():
            if self.use_eigen_decomp:
                self._round16('QA', 'dA')
                get_ops().transpose(self.state['QA'], self._E['A'])
            else:
                self._round16('A_inv')

    def _finish_G_inv(self, damping):
        """dGdA = 1/(dG dA^T + damping) on the rank that decomposed G (base.py:300-306); with a 16-bit
        inv_dtype everything the reference would store in it is rounded to it (base.py:435-441)."""
        if self._inv16():
            if self.use_eigen_decomp:
                self._round16('QG', 'dG')
                get_ops().transpose(self.state['QG'], self._E['G'])
            else:
                self._round16('G_inv')
        if self.prediv_eigenvalues and self.use_eigen_decomp:
            if 'dA' not in self.state:
                raise ValueError('compute_A_inv must be called before '
                                 'compute_G_inv if prediv_eigenvalues '
                                 'is True.')
            get_ops().outer_recip(self.state['dG'], self.state['dA'], damping, self.state['dGdA'])
            if self._inv16():
                self._round16('dGdA')

    def compute_A_inv(self, damping=0.001, ignore_rank=False):
        """Eigendecompose A on its assigned rank (kfac/layers/base.py:198-256)."""
        job = self._inv_job('A', ignore_rank)
        if job is not None:
            self._run_inv_job(job, damping)
            self._finish_A_inv()

    def compute_G_inv(self, damping=0.001, ignore_rank=False):
        """Eigendecompose G on its assigned rank and, with prediv_eigenvalues, build
        dGdA = 1/(dG dA^T + damping) there (kfac/layers/base.py:258-308)."""
        job = self._inv_job('G', ignore_rank)
        if job is not None:
            self._run_inv_job(job, damping)
            self._finish_G_inv(damping)

    def broadcast_inverses(self):
        """Ship QA, QG and dGdA (or dA, dG) over the inverse groups (base.py:129-171).
        Receivers must call ``finish_inverse_broadcast`` once the handles completed."""
        if not self.keep_inv_copy:
            return []
        if not self.use_eigen_decomp:
            return [comm.backend.broadcast(self._buf['A_inv'], src=self.compute_A_inv_rank,
                                           group=self.broadcast_A_inv_group),
                    comm.backend.broadcast(self._buf['G_inv'], src=self.compute_G_inv_rank,
                                           group=self.broadcast_G_inv_group)]
        ops = [comm.backend.broadcast(self._buf['QA'], src=self.compute_A_inv_rank,
                                      group=self.broadcast_A_inv_group),
               comm.backend.broadcast(self._buf['QG'], src=self.compute_G_inv_rank,
                                      group=self.broadcast_G_inv_group)]
        if self.prediv_eigenvalues:
            ops.append(comm.backend.broadcast(self._buf['dGdA'], src=self.compute_A_inv_rank,
                                              group=self.broadcast_A_inv_group))
        else:
            ops.append(comm.backend.broadcast(self.state['dA'], src=self.compute_A_inv_rank,
                                              group=self.broadcast_A_inv_group))
            ops.append(comm.backend.broadcast(self.state['dG'], src=self.compute_G_inv_rank,
                                              group=self.broadcast_G_inv_group))
        return ops

    def finish_inverse_broadcast(self):
        """Rebuild E = Q^T for the factors this rank received rather than computed."""
        if not self.keep_inv_copy or not self.use_eigen_decomp:
            return
        rank = comm.backend.rank()
        if rank != self.compute_A_inv_rank:
            get_ops().transpose(self.state['QA'], self._E['A'])
        if rank != self.compute_G_inv_rank:
            get_ops().transpose(self.state['QG'], self._E['G'])

    # ------------------------------------------------------------------ gradients
    def _get_bias_grad(self):
        return self._contig_grad(self.module.bias)

    def _get_weight_grad(self):
        return self._contig_grad(self.module.weight)

    @staticmethod
    def _contig_grad(p):
        g = p.grad
        if g is not None and not g.is_contiguous():
            g = g.contiguous()
            p.grad = g
        return g

    def get_gradient(self):
        """[dW.view(out,-1) | db]  (kfac/layers/base.py:310-319)."""
        g = self._get_weight_grad()
        g = g.view(g.size(0), -1)
        if self.has_bias:
            g = torch.cat([g, self._get_bias_grad().data.view(-1, 1)], 1)
        return g

    def _dims(self):
        return self.state['A'].shape[0], self.state['G'].shape[0]

    def _set_out(self, view):
        w = self.module.weight
        n_w = view.shape[1] - int(self.has_bias)
        self._out = view
        self.preconditioned_gradient = [view[:, :n_w].unflatten(1, tuple(w.shape[1:]))
                                        if w.dim() > 2 else view[:, :n_w]]
        if self.has_bias:
            self.preconditioned_gradient.append(view[:, n_w])

    def _ensure_out(self):
        if self._out is None:
            da, dg = self._weight_dims()
            buf, view = alloc2d(dg, da, self.module.weight.device, zero=True)
            self._buf['out'] = buf
            self._set_out(view)
        return self._out

    def _weight_dims(self):
        w = self.module.weight
        return w[0].numel() + int(self.has_bias), w.shape[0]

    def _precond_job(self, damping):
        """The persistent kfb_precond_job of this layer, with this step's gradient
        pointers (kfac/layers/base.py:321-362)."""
        if self.compute_grad_ranks is None:
            raise ValueError('Gradient preconditioning workers have not been '
                             'assigned yet. Have you called assign_workers() '
                             'yet?')
        if comm.backend.rank() not in self.compute_grad_ranks:
            return None
        if self._job is None:
            da, dg = self._dims()
            dev = self.state['A'].device
            out = self._ensure_out()
            bufs = {}
            names = (('gmat', (dg, da)), ('V', (dg, da)), ('T', (da, dg))) if self.use_eigen_decomp \
                else (('gmat', (dg, da)), ('T', (da, dg)))
            for name, (r, c) in names:
                self._buf[name], bufs[name] = alloc2d(r, c, dev, zero=True)
            if not self.use_eigen_decomp:
                # out = G_inv gmat A_inv  (kfac/layers/base.py:472-475): two GEMM stages
                self._job = PrecondJob(da=da, dg=dg, has_bias=self.has_bias, use_inverse=True,
                                       EA=self.state['A_inv'], QG=self.state['G_inv'],
                                       gmat=bufs['gmat'], out=out, T=bufs['T'])
            else:
                if 'dGdA' not in self.state:
                    self._set2d('dGdA', dg, da, dev)
                self._job = PrecondJob(da=da, dg=dg, has_bias=self.has_bias,
                                       EA=self._E['A'], QA=self.state['QA'],
                                       EG=self._E['G'], QG=self.state['QG'],
                                       dGdA=self.state['dGdA'], gmat=bufs['gmat'], V=bufs['V'],
                                       out=out, T=bufs['T'])
        if self.use_eigen_decomp and not self.prediv_eigenvalues:
            # 1/(dG dA^T + damping) with the CURRENT damping (base.py:468-469)
            get_ops().outer_recip(self.state['dG'], self.state['dA'], damping, self.state['dGdA'])
        self._job.weight_grad = self._get_weight_grad()
        self._job.bias_grad = self._get_bias_grad() if self.has_bias else None
        return self._job

    def compute_preconditioned_gradient(self, damping=0.001):
        job = self._precond_job(damping)
        if job is not None:
            get_ops().precondition([job], None)

    def broadcast_gradient(self):
        """Broadcast the preconditioned gradient inside this rank's gradient group
        (kfac/layers/base.py:173-196); receivers start from zeros.  One message per layer:
        the padded [dg x da] matrix carries weight and bias together."""
        if self.compute_grad_ranks is None:
            raise ValueError('Gradient compute ranks have not been assigned '
                             'yet. Use assign_workers().')
        self._ensure_out()
        src, group = self.broadcast_grad_groups[comm.backend.rank()]
        return [comm.backend.broadcast(self._buf['out'], src=src, group=group)]

    def _grad_item(self):
        if self._out is None:
            raise RuntimeError('self.compute_preconditioned_gradient() should'
                               ' be called before update_gradient()')
        return (self._get_weight_grad(), self._get_bias_grad() if self.has_bias else None,
                self._out)

    def update_gradient(self, scale=None):
        """Write (scale *) the preconditioned gradient into the module's .grad tensors, in
        place (kfac/layers/base.py:419-430)."""
        get_ops().scale_writeback([self._grad_item()], None, 0.0, None,
                                  1.0 if scale is None else float(scale))
