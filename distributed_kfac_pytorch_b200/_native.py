"""ctypes binding of ``libkfac_b200.so`` (C ABI in ``include/kfac_b200.h``).

PyTorch is used only for device memory and streams: every argument crosses the
boundary as a raw device pointer, a size or a scalar.  There is NO fallback: if the
shared library is missing, or a tensor is not on a CUDA device, the call raises.

The host logic talks to the kernels through the small ``ops`` object returned by
``get_ops()``; tests of the host logic swap in a CPU stand-in with ``set_ops()``
(tests/_cpu_ops.py, built on the oracle) so that placement, hooks, collectives and
state handling can be exercised without a GPU.
"""
import ctypes as C
import os

import torch

_LIB_NAME = 'libkfac_b200.so'
_lib = None
_ops = None

KFB_SEG_ROWS, KFB_SEG_NCHW, KFB_SEG_IM2COL = 0, 1, 2
KFB_DTYPES = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


class NativeError(RuntimeError):
    pass


# ---------------------------------------------------------------------------- structs
class CSegment(C.Structure):
    _fields_ = [('kind', C.c_int32), ('has_bias', C.c_int32), ('data', C.c_void_p),
                ('n', C.c_int64), ('c', C.c_int64), ('h', C.c_int64), ('w', C.c_int64),
                ('kh', C.c_int32), ('kw', C.c_int32), ('sh', C.c_int32), ('sw', C.c_int32),
                ('ph', C.c_int32), ('pw', C.c_int32), ('ld', C.c_int64),
                ('weight', C.c_float), ('dtype', C.c_int32), ('dev_scale', C.c_void_p)]


class CFactorJob(C.Structure):
    _fields_ = [('d', C.c_int32), ('n_segments', C.c_int32), ('segments', C.POINTER(CSegment)),
                ('state', C.c_void_p), ('ld_state', C.c_int64), ('seed_identity', C.c_int32),
                ('write_state', C.c_int32), ('tri_out', C.c_void_p), ('dev_rescale', C.c_void_p)]


class CFilterJob(C.Structure):
    _fields_ = [('data', C.c_void_p), ('numel', C.c_int64), ('dtype', C.c_int32), ('group', C.c_int32),
                ('rows', C.c_float), ('host_inv_scale_sq', C.c_float), ('loss_scale', C.c_void_p),
                ('dev_scale', C.c_void_p)]


class CFilterGroup(C.Structure):
    _fields_ = [('dev_rescale', C.c_void_p), ('renormalise', C.c_int32), ('reserved', C.c_int32)]


class CTriJob(C.Structure):
    _fields_ = [('tri', C.c_void_p), ('state', C.c_void_p), ('ld_state', C.c_int64),
                ('d', C.c_int32), ('reserved', C.c_int32)]


class CSyevdJob(C.Structure):
    _fields_ = [('factor', C.c_void_p), ('ld_factor', C.c_int64), ('E', C.c_void_p), ('ld_e', C.c_int64),
                ('Q', C.c_void_p), ('ld_q', C.c_int64), ('w', C.c_void_p), ('dev_info', C.c_void_p),
                ('d', C.c_int32), ('reserved', C.c_int32)]


class CPrecondJob(C.Structure):
    _fields_ = [('da', C.c_int32), ('dg', C.c_int32), ('has_bias', C.c_int32),
                ('use_inverse', C.c_int32), ('weight_grad', C.c_void_p), ('bias_grad', C.c_void_p),
                ('EA', C.c_void_p), ('QA', C.c_void_p), ('ld_a', C.c_int64),
                ('EG', C.c_void_p), ('QG', C.c_void_p), ('ld_g', C.c_int64),
                ('dGdA', C.c_void_p), ('ld_dgda', C.c_int64),
                ('gmat', C.c_void_p), ('V', C.c_void_p), ('out', C.c_void_p), ('ld_v', C.c_int64),
                ('T', C.c_void_p), ('ld_t', C.c_int64)]


class CGradJob(C.Structure):
    _fields_ = [('weight_grad', C.c_void_p), ('bias_grad', C.c_void_p), ('v', C.c_void_p),
                ('ld_v', C.c_int64), ('da', C.c_int32), ('dg', C.c_int32),
                ('has_bias', C.c_int32), ('reserved', C.c_int32)]


_SIGNATURES = {
    'kfb_version': (C.c_int, []),
    'kfb_last_error': (C.c_char_p, []),
    'kfb_init': (C.c_int, [C.c_int]),
    'kfb_set_mma_mode': (C.c_int, [C.c_int]),
    'kfb_get_mma_mode': (C.c_int, []),
    'kfb_launch_count': (C.c_ulonglong, []),
    'kfb_factor_workspace_bytes': (C.c_size_t, [C.POINTER(CFactorJob), C.c_int32]),
    'kfb_factor_workspace_dedicate': (C.c_int, [C.c_void_p, C.c_size_t]),
    'kfb_factor_workspace_release': (C.c_int, [C.c_void_p]),
    'kfb_factor_update': (C.c_int, [C.POINTER(CFactorJob), C.c_int32, C.c_float, C.c_void_p,
                                    C.c_size_t, C.c_void_p]),
    'kfb_capture_filter': (C.c_int, [C.POINTER(CFilterJob), C.c_int32, C.POINTER(CFilterGroup), C.c_int32, C.c_void_p]),
    'kfb_narrow': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    'kfb_widen': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    'kfb_static_scratch_bytes': (C.c_size_t, []),
    'kfb_factor_update_static': (C.c_int, [C.POINTER(CFactorJob), C.c_int32, C.c_float, C.c_void_p]),
    'kfb_tri_unpack': (C.c_int, [C.POINTER(CTriJob), C.c_int32, C.c_float, C.c_void_p]),
    'kfb_tri_pack': (C.c_int, [C.POINTER(CTriJob), C.c_int32, C.c_void_p]),
    'kfb_syevd_workspace_bytes': (C.c_size_t, [C.c_int32]),
    'kfb_syevd': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p,
                            C.c_int64, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    'kfb_syevd_batch_workspace_bytes': (C.c_size_t, [C.POINTER(CSyevdJob), C.c_int32]),
    'kfb_syevd_batch': (C.c_int, [C.POINTER(CSyevdJob), C.c_int32, C.c_void_p, C.c_size_t, C.c_void_p]),
    'kfb_syevd_lane_plan': (C.c_int, [C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32),
                                      C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_double)]),
    'kfb_cholesky_inverse_workspace_bytes': (C.c_size_t, [C.c_int32]),
    'kfb_cholesky_inverse': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_void_p, C.c_int64,
                                       C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    'kfb_transpose': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                C.c_int64, C.c_void_p]),
    'kfb_outer_recip': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_float,
                                  C.c_void_p, C.c_int64, C.c_void_p]),
    'kfb_precondition': (C.c_int, [C.POINTER(CPrecondJob), C.c_int32, C.c_void_p, C.c_void_p]),
    'kfb_gemm_tn': (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                              C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    'kfb_dot_accumulate': (C.c_int, [C.POINTER(CGradJob), C.c_int32, C.c_void_p, C.c_void_p]),
    'kfb_scale_writeback': (C.c_int, [C.POINTER(CGradJob), C.c_int32, C.c_void_p, C.c_float,
                                      C.c_float, C.c_float, C.c_void_p]),
}

EXPORTED_SYMBOLS = tuple(sorted(_SIGNATURES))


def library_path():
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), _LIB_NAME)


def load_library():
    """dlopen the extension and declare every prototype.  Raises NativeError if the
    file is missing (run ``python -c 'import __graft_entry__ as g; g.build()'``)."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise NativeError('%s not found: the CUDA extension has not been built '
                          '(python -c "import __graft_entry__ as g; g.build()"). '
                          'There is no CPU fallback.' % path)
    lib = C.CDLL(path)
    for name, (restype, argtypes) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


# ------------------------------------------------------------------------ job objects
class Segment(object):
    """One operand of a factor: a captured activation / grad_output tensor plus the
    layout it is read through (see kfb_segment)."""
    __slots__ = ('kind', 'has_bias', 'tensor', 'n', 'c', 'h', 'w', 'kernel', 'stride',
                 'padding', 'ld', 'weight', 'dev_scale')

    def __init__(self, kind, tensor, n, c, weight, has_bias=False, h=1, w=1, kernel=(1, 1),
                 stride=(1, 1), padding=(0, 0), ld=0):
        self.kind, self.tensor, self.n, self.c, self.weight = kind, tensor, n, c, weight
        self.has_bias, self.h, self.w = has_bias, h, w
        self.kernel, self.stride, self.padding, self.ld = kernel, stride, padding, ld
        self.dev_scale = None      # 1-element device tensor multiplying `weight` (0 drops the segment)


class FactorJob(object):
    __slots__ = ('d', 'segments', 'state', 'seed_identity', 'write_state', 'tri_out', 'dev_rescale', 'filter')

    def __init__(self, d, segments, state, seed_identity, write_state=True, tri_out=None):
        self.d, self.segments, self.state = d, segments, state
        self.seed_identity, self.write_state, self.tri_out = seed_identity, write_state, tri_out
        self.dev_rescale = None    # 1-element device tensor: >= 0 multiplies F, < 0 skips the update
        self.filter = None         # CaptureFilter that fills dev_rescale and the segments' dev_scale


class CaptureFilter(object):
    """AMP bookkeeping of one G factor: its captures, their loss scale, and the device scalars the
    filter kernel writes (one per capture + one per factor)."""
    __slots__ = ('tensors', 'rows', 'scales', 'renormalise', 'dev_scales', 'dev_rescale')

    def __init__(self, tensors, rows, scales, renormalise, dev_scales, dev_rescale):
        self.tensors, self.rows, self.scales, self.renormalise = tensors, rows, scales, renormalise
        self.dev_scales, self.dev_rescale = dev_scales, dev_rescale


class PrecondJob(object):
    """Persistent per-layer description of the precondition GEMM chain; only
    ``weight_grad`` / ``bias_grad`` change from step to step."""
    __slots__ = ('da', 'dg', 'has_bias', 'use_inverse', 'weight_grad', 'bias_grad', 'EA', 'QA', 'EG', 'QG',
                 'dGdA', 'gmat', 'V', 'out', 'T', '_c')

    def __init__(self, **kw):
        self._c = None
        self.weight_grad = self.bias_grad = None
        self.use_inverse = False       # True: EA holds A_inv, QG holds G_inv, out = G_inv gmat A_inv
        self.QA = self.EG = self.dGdA = self.V = None
        for k, v in kw.items():
            setattr(self, k, v)


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _check_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NativeError('K-FAC B200 kernels need CUDA tensors (got a %s tensor); there is '
                              'no CPU fallback' % t.device)


def _check_f32(*tensors):
    for t in tensors:
        if t is not None and t.dtype != torch.float32:
            raise NativeError('kernels take float32 tensors, got %s' % t.dtype)


def _pitch(t):
    """Row pitch of a 2-D tensor whose rows are contiguous."""
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1):
        raise NativeError('expected a row-major 2-D tensor, got shape %s stride %s'
                          % (tuple(t.shape), t.stride()))
    return t.stride(0) if t.shape[0] > 1 else max(t.stride(0), t.shape[1])


class CudaOps(object):
    """The product path: every method is one call into libkfac_b200.so on the current
    CUDA stream of the tensors' device."""

    name = 'cuda'

    def __init__(self):
        self.lib = load_library()
        self._inited = set()
        self._workspaces = {}
        self._factor_ws = {}
        self._keepalive = None

    # -- helpers
    on_native_call = None      # profiling hook: called with the entry point's name right before the call

    def _call(self, name, *args):
        if self.on_native_call is not None:
            self.on_native_call(name)
        rc = getattr(self.lib, name)(*args)
        if rc != 0:
            raise NativeError('%s failed (%d): %s' % (name, rc, self.lib.kfb_last_error().decode()))

    def _enter(self, tensor):
        _check_cuda(tensor)
        dev = tensor.device.index if tensor.device.index is not None else torch.cuda.current_device()
        if torch.cuda.current_device() != dev:
            torch.cuda.set_device(dev)
        self._call('kfb_init', dev)     # also selects the per-thread context; cheap when warm
        return torch.cuda.current_stream(dev).cuda_stream

    def workspace(self, nbytes, device):
        key = (device.type, device.index)
        buf = self._workspaces.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
            self._workspaces[key] = buf
        return buf

    def factor_workspace(self, nbytes, device):
        """The factor update's own workspace: nothing else writes into it, and the library is told so
        (kfb_factor_workspace_dedicate), which lets it keep its split-K accumulators zeroed in place."""
        key = (device.type, device.index)
        buf = self._factor_ws.get(key)
        if buf is None or buf.numel() < nbytes:
            if buf is not None:
                self._call('kfb_factor_workspace_release', buf.data_ptr())
            buf = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
            self._call('kfb_factor_workspace_dedicate', buf.data_ptr(), buf.numel())
            self._factor_ws[key] = buf
        return buf

    def __del__(self):
        try:
            for buf in self._factor_ws.values():
                self.lib.kfb_factor_workspace_release(buf.data_ptr())
        except Exception:
            pass

    def set_mma_mode(self, mode):
        self._call('kfb_set_mma_mode', int(mode))

    def get_mma_mode(self):
        return self.lib.kfb_get_mma_mode()

    def launch_count(self):
        return int(self.lib.kfb_launch_count())

    # -- factors
    def _factor_cjobs(self, jobs):
        n = len(jobs)
        cjobs = (CFactorJob * n)()
        keep = []
        for cj, j in zip(cjobs, jobs):
            segs = (CSegment * len(j.segments))()
            for cs, s in zip(segs, j.segments):
                _check_cuda(s.tensor)
                if s.tensor.dtype not in KFB_DTYPES:
                    raise NativeError('captured tensors must be float32, float16 or bfloat16, got %s'
                                      % s.tensor.dtype)
                cs.dtype = KFB_DTYPES[s.tensor.dtype]
                if not s.tensor.is_contiguous() and s.kind != KFB_SEG_ROWS:
                    raise NativeError('captured conv tensors must be contiguous NCHW')
                cs.kind, cs.has_bias, cs.data = s.kind, int(s.has_bias), s.tensor.data_ptr()
                cs.n, cs.c, cs.h, cs.w = s.n, s.c, s.h, s.w
                cs.kh, cs.kw = s.kernel
                cs.sh, cs.sw = s.stride
                cs.ph, cs.pw = s.padding
                cs.ld, cs.weight = s.ld, s.weight
                cs.dev_scale = _ptr(s.dev_scale)
            keep.append(segs)
            _check_cuda(j.state, j.tri_out)
            _check_f32(j.state, j.tri_out)
            cj.d, cj.n_segments, cj.segments = j.d, len(j.segments), segs
            cj.state, cj.ld_state = j.state.data_ptr(), _pitch(j.state)
            cj.seed_identity, cj.write_state = int(j.seed_identity), int(j.write_state)
            cj.tri_out = _ptr(j.tri_out)
            cj.dev_rescale = _ptr(j.dev_rescale)
        return cjobs, keep

    def capture_filter(self, filters):
        """filters: list of CaptureFilter.  One native call decides, on the device, which captures are
        finite, folds 1 / loss_scale^2 into their weights and renormalises the factors that lost some."""
        if not filters:
            return
        stream = self._enter(filters[0].tensors[0])
        n = sum(len(f.tensors) for f in filters)
        cjobs = (CFilterJob * n)()
        cgroups = (CFilterGroup * len(filters))()
        i = 0
        for g, f in enumerate(filters):
            cgroups[g].dev_rescale, cgroups[g].renormalise = f.dev_rescale.data_ptr(), int(f.renormalise)
            for k, (t, rows, scale) in enumerate(zip(f.tensors, f.rows, f.scales)):
                _check_cuda(t)
                if not t.is_contiguous():
                    raise NativeError('captured tensors must be contiguous')
                cj = cjobs[i]
                cj.data, cj.numel, cj.dtype, cj.group = t.data_ptr(), t.numel(), KFB_DTYPES[t.dtype], g
                cj.rows = float(rows)
                if torch.is_tensor(scale):
                    _check_cuda(scale)
                    _check_f32(scale)
                    cj.loss_scale, cj.host_inv_scale_sq = scale.data_ptr(), 0.0
                else:
                    cj.loss_scale, cj.host_inv_scale_sq = None, 1.0 / (float(scale) * float(scale))
                cj.dev_scale = f.dev_scales[k:k + 1].data_ptr()
                i += 1
        self._call('kfb_capture_filter', cjobs, n, cgroups, len(filters), stream)

    def narrow(self, src, dst16):
        """dst16 (float16 / bfloat16, flat) <- round(src fp32, flat)."""
        stream = self._enter(src)
        _check_cuda(dst16)
        _check_f32(src)
        if not (src.is_contiguous() and dst16.is_contiguous() and src.numel() == dst16.numel()):
            raise NativeError('narrow: contiguous tensors of equal size expected')
        self._call('kfb_narrow', src.data_ptr(), dst16.data_ptr(), src.numel(), KFB_DTYPES[dst16.dtype], stream)

    def widen(self, src16, dst):
        stream = self._enter(dst)
        _check_cuda(src16)
        _check_f32(dst)
        if not (dst.is_contiguous() and src16.is_contiguous() and dst.numel() == src16.numel()):
            raise NativeError('widen: contiguous tensors of equal size expected')
        self._call('kfb_widen', src16.data_ptr(), dst.data_ptr(), dst.numel(), KFB_DTYPES[src16.dtype], stream)

    def factor_update(self, jobs, alpha):
        if not jobs:
            return
        stream = self._enter(jobs[0].state)
        self.capture_filter([j.filter for j in jobs if j.filter is not None])
        cjobs, keep = self._factor_cjobs(jobs)
        n = len(jobs)
        nbytes = self.lib.kfb_factor_workspace_bytes(cjobs, n)
        ws = self.factor_workspace(nbytes, jobs[0].state.device)
        self._call('kfb_factor_update', cjobs, n, float(alpha), ws.data_ptr(), ws.numel(), stream)

    def factor_update_static(self, jobs, alpha):
        """kfb_factor_update_static: same result as factor_update, but the scratch is the
        library's built-in fixed array, so nothing is allocated (not even by torch).  Raises
        NativeError if the jobs need more than ``static_scratch_bytes()``."""
        if not jobs:
            return
        stream = self._enter(jobs[0].state)
        cjobs, keep = self._factor_cjobs(jobs)
        self._call('kfb_factor_update_static', cjobs, len(jobs), float(alpha), stream)

    def static_scratch_bytes(self):
        return int(self.lib.kfb_static_scratch_bytes())

    def _tri_jobs(self, pairs):
        cjobs = (CTriJob * len(pairs))()
        for cj, (tri, state) in zip(cjobs, pairs):
            _check_cuda(tri, state)
            _check_f32(tri, state)
            d = state.shape[0]
            if tri.numel() != d * (d + 1) // 2 or not tri.is_contiguous():
                raise NativeError('packed triangle has the wrong size')
            cj.tri, cj.state, cj.ld_state, cj.d = tri.data_ptr(), state.data_ptr(), _pitch(state), d
        return cjobs

    def tri_unpack(self, pairs, scale=1.0):
        if pairs:
            stream = self._enter(pairs[0][1])
            self._call('kfb_tri_unpack', self._tri_jobs(pairs), len(pairs), float(scale), stream)

    def tri_pack(self, pairs):
        if pairs:
            stream = self._enter(pairs[0][1])
            self._call('kfb_tri_pack', self._tri_jobs(pairs), len(pairs), stream)

    # -- inverses
    def syevd(self, factor, E, Q, w, info):
        stream = self._enter(factor)
        _check_cuda(E, Q, w, info)
        _check_f32(factor, E, Q, w)
        d = factor.shape[0]
        nbytes = self.lib.kfb_syevd_workspace_bytes(d)
        ws = self.workspace(nbytes, factor.device)
        self._call('kfb_syevd', factor.data_ptr(), _pitch(factor), d, E.data_ptr(), _pitch(E),
                   Q.data_ptr(), _pitch(Q), w.data_ptr(), ws.data_ptr(), ws.numel(),
                   info.data_ptr(), stream)

    def syevd_batch(self, jobs):
        """jobs: list of (factor, E, Q, w, info) tensors; all decompositions run concurrently on
        the library's worker lanes and are ordered after / before the current stream."""
        if not jobs:
            return
        stream = self._enter(jobs[0][0])
        cjobs = (CSyevdJob * len(jobs))()
        for cj, (factor, E, Q, w, info) in zip(cjobs, jobs):
            _check_cuda(factor, E, Q, w, info)
            _check_f32(factor, E, Q, w)
            cj.factor, cj.ld_factor = factor.data_ptr(), _pitch(factor)
            cj.E, cj.ld_e, cj.Q, cj.ld_q = E.data_ptr(), _pitch(E), Q.data_ptr(), _pitch(Q)
            cj.w, cj.dev_info, cj.d = w.data_ptr(), info.data_ptr(), factor.shape[0]
        nbytes = self.lib.kfb_syevd_batch_workspace_bytes(cjobs, len(jobs))
        ws = self.workspace(nbytes, jobs[0][0].device)
        self._call('kfb_syevd_batch', cjobs, len(jobs), ws.data_ptr(), ws.numel(), stream)

    def cholesky_inverse(self, factor, damping, out, info):
        """out = (factor + damping I)^-1 (full symmetric), cuSOLVER potrf + potri."""
        stream = self._enter(factor)
        _check_cuda(out, info)
        _check_f32(factor, out)
        d = factor.shape[0]
        ws = self.workspace(self.lib.kfb_cholesky_inverse_workspace_bytes(d), factor.device)
        self._call('kfb_cholesky_inverse', factor.data_ptr(), _pitch(factor), d, float(damping), out.data_ptr(),
                   _pitch(out), ws.data_ptr(), ws.numel(), info.data_ptr(), stream)

    def transpose(self, src, dst):
        stream = self._enter(src)
        _check_cuda(dst)
        _check_f32(src, dst)
        self._call('kfb_transpose', src.data_ptr(), _pitch(src), src.shape[0], src.shape[1],
                   dst.data_ptr(), _pitch(dst), stream)

    def outer_recip(self, dG, dA, damping, out):
        stream = self._enter(out)
        _check_cuda(dG, dA)
        _check_f32(dG, dA, out)
        self._call('kfb_outer_recip', dG.data_ptr(), dA.data_ptr(), dG.numel(), dA.numel(),
                   float(damping), out.data_ptr(), _pitch(out), stream)

    # -- preconditioning
    @staticmethod
    def _precond_c(j):
        c = j._c
        if c is None:
            c = CPrecondJob()
            _check_cuda(j.EA, j.QA, j.EG, j.QG, j.dGdA, j.gmat, j.V, j.out, j.T)
            _check_f32(j.EA, j.QA, j.EG, j.QG, j.dGdA, j.gmat, j.V, j.out, j.T)
            c.da, c.dg, c.has_bias, c.use_inverse = j.da, j.dg, int(j.has_bias), int(j.use_inverse)
            c.EA, c.QA, c.ld_a = j.EA.data_ptr(), _ptr(j.QA), _pitch(j.EA)
            c.EG, c.QG, c.ld_g = _ptr(j.EG), j.QG.data_ptr(), _pitch(j.QG)
            if not j.use_inverse:
                assert _pitch(j.QA) == c.ld_a and _pitch(j.EG) == c.ld_g
                c.dGdA, c.ld_dgda = j.dGdA.data_ptr(), _pitch(j.dGdA)
                assert _pitch(j.V) == _pitch(j.out)
            c.gmat, c.V, c.out, c.ld_v = j.gmat.data_ptr(), _ptr(j.V), j.out.data_ptr(), _pitch(j.out)
            assert _pitch(j.gmat) == c.ld_v
            c.T, c.ld_t = j.T.data_ptr(), _pitch(j.T)
            j._c = c
        wg, bg = j.weight_grad, j.bias_grad
        _check_cuda(wg, bg)
        _check_f32(wg, bg)
        if not wg.is_contiguous() or (bg is not None and not bg.is_contiguous()):
            raise NativeError('parameter gradients must be contiguous')
        c.weight_grad, c.bias_grad = wg.data_ptr(), _ptr(bg)
        return c

    def precondition(self, jobs, vg_acc):
        if not jobs:
            return
        stream = self._enter(jobs[0].out)
        cjobs = (CPrecondJob * len(jobs))(*[self._precond_c(j) for j in jobs])
        self._call('kfb_precondition', cjobs, len(jobs), _ptr(vg_acc), stream)

    def gemm_tn(self, A, B, D, mul=None):
        """D = A @ B.T (optionally * mul), all row-major 2-D fp32 views."""
        stream = self._enter(D)
        _check_cuda(A, B, mul)
        _check_f32(A, B, D, mul)
        self._call('kfb_gemm_tn', A.data_ptr(), _pitch(A), B.data_ptr(), _pitch(B), D.data_ptr(), _pitch(D),
                   A.shape[0], B.shape[0], A.shape[1], _ptr(mul), 0 if mul is None else _pitch(mul), stream)

    @staticmethod
    def _grad_jobs(items):
        """items: (weight_grad, bias_grad|None, v [dg x da view])"""
        cjobs = (CGradJob * len(items))()
        for cj, (wg, bg, v) in zip(cjobs, items):
            _check_cuda(wg, bg, v)
            _check_f32(wg, bg, v)
            if not wg.is_contiguous() or (bg is not None and not bg.is_contiguous()):
                raise NativeError('parameter gradients must be contiguous')
            cj.weight_grad, cj.bias_grad, cj.v, cj.ld_v = wg.data_ptr(), _ptr(bg), v.data_ptr(), _pitch(v)
            cj.dg, cj.da, cj.has_bias = v.shape[0], v.shape[1], int(bg is not None)
        return cjobs

    def dot_accumulate(self, items, vg_acc):
        if items:
            stream = self._enter(items[0][2])
            self._call('kfb_dot_accumulate', self._grad_jobs(items), len(items), vg_acc.data_ptr(), stream)

    def scale_writeback(self, items, vg_acc, lr, kl_clip, fixed_scale=1.0):
        if items:
            stream = self._enter(items[0][2])
            self._call('kfb_scale_writeback', self._grad_jobs(items), len(items), _ptr(vg_acc),
                       float(lr), float(kl_clip if kl_clip is not None else 0.0),
                       float(fixed_scale), stream)


def syevd_lane_plan(dims, num_sms=148, max_lanes=6):
    """kfb_syevd_lane_plan: (lane of every factor, SMs per lane, modelled ms).  Host only."""
    lib = load_library()
    n = len(dims)
    cd = (C.c_int32 * n)(*dims)
    lane_of = (C.c_int32 * n)()
    lane_sms = (C.c_int32 * 8)()
    n_lanes, ms = C.c_int32(), C.c_double()
    rc = lib.kfb_syevd_lane_plan(cd, n, num_sms, max_lanes, lane_of, lane_sms, C.byref(n_lanes), C.byref(ms))
    if rc != 0:
        raise NativeError('kfb_syevd_lane_plan failed (%d): %s' % (rc, lib.kfb_last_error().decode()))
    return list(lane_of), list(lane_sms)[:n_lanes.value], ms.value


def get_ops():
    """The ops object used by the host logic.  Loads the CUDA library on first use and
    fails loudly if it is not there."""
    global _ops
    if _ops is None:
        _ops = CudaOps()
    return _ops


def set_ops(ops):
    """Install another ops implementation (tests only).  ``None`` restores the default."""
    global _ops
    _ops = ops
