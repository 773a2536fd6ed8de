"""Parity of every C-ABI entry point against the oracle on a B200 (-m gpu).

Each test builds jobs on the GPU, runs them through libkfac_b200.so (CudaOps), builds the
same jobs on CPU copies of the tensors and runs them through the oracle-backed stand-in
(tests/_cpu_ops.py), and compares.  Tolerances are north_star's: factors rtol 1e-3 (measured
against the factor's own scale: |err| <= 1e-3 * sqrt(F_ii F_jj), since elementwise rtol is
meaningless for entries that cancel to ~0), preconditioned gradients rtol 1e-2 / atol 1e-4."""
import math

import pytest
import torch

from conftest import load_golden
from distributed_kfac_pytorch_b200 import _native as N
from distributed_kfac_pytorch_b200.layers.base import alloc2d
from _cpu_ops import OracleOps, triu_pack
from oracle import kfac_oracle as O

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


@pytest.fixture(params=[0, 1], ids=['simt', 'tcgen05'])
def ops(request):
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    o.set_mma_mode(request.param)
    yield o
    o.set_mma_mode(1)


def assert_factor_close(got, ref, tol=1e-3):
    got = got.detach().cpu().double()
    ref = ref.detach().cpu().double()
    dscale = ref.diagonal().clamp_min(0).sqrt()
    bound = tol * torch.outer(dscale, dscale) + 1e-9
    err = (got - ref).abs()
    worst = (err / bound).max().item()
    assert worst <= 1.0, 'factor error %.3g x tolerance' % worst
    assert torch.equal(got, got.t()), 'factor is not exactly symmetric'


def seg_copy(s, device):
    t = s.tensor.to(device)
    return N.Segment(s.kind, t, s.n, s.c, s.weight, has_bias=s.has_bias, h=s.h, w=s.w,
                     kernel=s.kernel, stride=s.stride, padding=s.padding, ld=s.ld)


def run_both(ops, d, segments, alpha=0.95, old=None, with_tri=True):
    """Run one factor job on GPU and on the oracle; returns (gpu_state, gpu_tri, ref_state)."""
    buf, state = alloc2d(d, d, DEV)
    buf.fill_(float('nan'))
    if old is not None:
        state.copy_(old)
    tri = torch.full((d * (d + 1) // 2,), float('nan'), device=DEV) if with_tri else None
    ops.factor_update([N.FactorJob(d, segments, state, old is None, True, tri)], alpha)
    torch.cuda.synchronize()
    ref_state = torch.empty(d, d) if old is None else old.clone()
    OracleOps().factor_update([N.FactorJob(d, [seg_copy(s, 'cpu') for s in segments], ref_state,
                                           old is None, True, None)], alpha)
    return state, tri, ref_state


def conv_segments(x, k, s, p, has_bias, dev=DEV):
    n, c, h, w = x.shape
    oh = (h + 2 * p[0] - k[0]) // s[0] + 1
    ow = (w + 2 * p[1] - k[1]) // s[1] + 1
    S = oh * ow
    seg = N.Segment(N.KFB_SEG_IM2COL, x.to(dev).contiguous(), n, c, 1.0 / (S * S * n * S),
                    has_bias=has_bias, h=h, w=w, kernel=k, stride=s, padding=p)
    return c * k[0] * k[1] + int(has_bias), [seg]


CONV_CASES = [
    # (N, C, H, W, kernel, stride, padding, bias)
    (4, 3, 17, 13, (3, 3), (1, 1), (1, 1), False),
    (3, 5, 16, 16, (3, 3), (2, 2), (1, 1), True),
    (2, 3, 31, 29, (7, 7), (2, 2), (3, 3), False),     # conv1-like, d = 147
    (5, 16, 9, 9, (1, 1), (2, 2), (0, 0), False),      # downsample 1x1 stride 2
    (2, 40, 7, 7, (3, 3), (1, 1), (1, 1), False),      # d = 360, 7x7 map (49 not /4)
    (3, 8, 10, 12, (3, 5), (2, 1), (0, 2), True),      # asymmetric everything
    (1, 1, 3, 3, (3, 3), (1, 1), (0, 0), False),       # single output pixel, K = 1
    (2, 70, 14, 14, (3, 3), (1, 1), (1, 1), False),    # d = 630 > several tiles
    (9, 128, 28, 28, (3, 3), (1, 1), (1, 1), False),   # d = 1152: whole tiles only, K chunks
    (33, 64, 14, 14, (3, 3), (2, 2), (1, 1), True),    # d = 577: bias in a tile of its own
    (20, 256, 16, 16, (1, 1), (2, 2), (0, 0), False),  # sub-sampled 1x1 -> TMA-fed NCHW copy
    (3, 64, 12, 12, (1, 1), (2, 2), (0, 0), True),     # sub-sampled 1x1 with bias
]


@pytest.mark.parametrize('case', CONV_CASES, ids=lambda c: 'x'.join(map(str, c[:4])) + '_k%d' % c[4][0])
def test_conv_A_factor(ops, case):
    n, c, h, w, k, s, p, bias = case
    g = torch.Generator().manual_seed(hash(case) % 1000)
    x = torch.randn(n, c, h, w, generator=g)
    d, segs = conv_segments(x, k, s, p, bias)
    state, tri, ref = run_both(ops, d, segs)
    assert_factor_close(state, ref)
    torch.testing.assert_close(tri.cpu(), triu_pack(state.cpu()), rtol=0, atol=0)
    # closed form against the oracle's own conv2d_A (not via the stand-in)
    direct = O.factor_ema(O.conv2d_A([x], k, s, p, bias), None, 0.95)
    assert_factor_close(state, direct)


@pytest.mark.parametrize('shape', [(4, 6, 5, 5), (32, 64, 7, 7), (3, 130, 14, 14), (2, 16, 56, 56), (1, 3, 1, 1),
                                   (5, 300, 6, 6), (40, 70, 8, 8), (24, 260, 14, 14)])
def test_conv_G_factor_nchw(ops, shape):
    n, c, h, w = shape
    g = torch.randn(*shape, generator=torch.Generator().manual_seed(7))
    S = h * w
    seg = N.Segment(N.KFB_SEG_NCHW, g.to(DEV), n, c, 1.0 / (S * S * n * S), h=h, w=w)
    state, tri, ref = run_both(ops, c, [seg])
    assert_factor_close(state, ref)
    assert_factor_close(state, O.factor_ema(O.conv2d_G([g]), None, 0.95))


@pytest.mark.parametrize('case', [c for c in CONV_CASES if c[0] * c[1] >= 15], ids=lambda c: 'x'.join(map(str, c[:4])) + '_k%d' % c[4][0])
def test_conv_A_staged_and_in_place_agree(monkeypatch, case):
    """The tensor-core SYRK reads im2col operands from zero-padded fp32 staging copies (1x1 strided
    convs: dense sub-sampled copies); KFB_SYRK_STAGE=0 gathers from the capture in place with border
    masks.  Same products, so the two agree far inside the oracle tolerance - for fp32 and 16-bit
    captures."""
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    o.set_mma_mode(1)
    n, c, h, w, k, s, p, bias = case
    for dtype in (torch.float32, torch.bfloat16):
        x = torch.randn(n, c, h, w, generator=torch.Generator().manual_seed(5)).to(dtype)
        ref = O.factor_ema(O.conv2d_A([x.float()], k, s, p, bias), None, 0.95)
        out = []
        for env in ('1', '0'):
            monkeypatch.setenv('KFB_SYRK_STAGE', env)
            d, segs = conv_segments(x, k, s, p, bias)
            buf, state = alloc2d(d, d, DEV)
            o.factor_update([N.FactorJob(d, segs, state, True, True, None)], 0.95)
            torch.cuda.synchronize()
            assert_factor_close(state, ref)
            out.append(state.clone())
        assert_factor_close(out[0], out[1], tol=1e-5)


def test_cta_pair_syrk_matches_single_cta(monkeypatch):
    """KFB_SYRK_2CTA=<tiles> sends factors of at least that many 128-wide tiles per side to the
    cta_group::2 build of the SYRK (leader issues, peer relays its landed stages, accumulators
    double-buffered).  Off by default (measured slower while the gathers are load/store-unit bound);
    kept parity-tested: odd tile counts, diagonal-only factors, bias tiles, TMA-fed and gathered
    operands, fused and chunked jobs."""
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    o.set_mma_mode(1)
    gen = torch.Generator().manual_seed(33)
    cases = [(9, 128, 28, 28, (3, 3), (1, 1), (1, 1), False), (33, 64, 14, 14, (3, 3), (2, 2), (1, 1), True),
             (2, 70, 14, 14, (3, 3), (1, 1), (1, 1), False), (2, 3, 31, 29, (7, 7), (2, 2), (3, 3), False)]
    for n, c, h, w, k, s, p, bias in cases:
        x = torch.randn(n, c, h, w, generator=gen)
        ref = O.factor_ema(O.conv2d_A([x], k, s, p, bias), None, 0.95)
        out = []
        for env in ('1', '0'):
            monkeypatch.setenv('KFB_SYRK_2CTA', env)
            d, segs = conv_segments(x, k, s, p, bias)
            buf, state = alloc2d(d, d, DEV)
            tri = torch.full((d * (d + 1) // 2,), float('nan'), device=DEV)
            o.factor_update([N.FactorJob(d, segs, state, True, True, tri)], 0.95)
            torch.cuda.synchronize()
            assert_factor_close(state, ref)
            torch.testing.assert_close(tri.cpu(), triu_pack(state.cpu()), rtol=0, atol=0)
            out.append(state.clone())
        assert_factor_close(out[0], out[1], tol=1e-5)
    for shape in [(24, 260, 14, 14), (3, 300, 7, 7), (2, 16, 56, 56)]:      # NCHW: TMA-fed / gathered
        n, c, h, w = shape
        g = torch.randn(*shape, generator=gen)
        S = h * w
        ref = O.factor_ema(O.conv2d_G([g]), None, 0.95)
        for env in ('1', '0'):
            monkeypatch.setenv('KFB_SYRK_2CTA', env)
            seg = N.Segment(N.KFB_SEG_NCHW, g.to(DEV), n, c, 1.0 / (S * S * n * S), h=h, w=w)
            buf, state = alloc2d(c, c, DEV)
            o.factor_update([N.FactorJob(c, [seg], state, True, True, None)], 0.95)
            torch.cuda.synchronize()
            assert_factor_close(state, ref)


def test_stream_k_schedule_matches_default(monkeypatch):
    """KFB_SYRK_STREAMK=1 cuts the SYRK work into one contiguous share per CTA (pieces of arbitrary K
    length, accumulators kept in TMEM across a piece) instead of fixed K chunks handed out round-robin.
    Off by default (measured slower on ResNet-50); a batch of jobs large enough to be cut inside units
    must still give the oracle's factors."""
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    o.set_mma_mode(1)
    gen = torch.Generator().manual_seed(34)
    xs = [torch.randn(9, 128, 28, 28, generator=gen), torch.randn(6, 70, 14, 14, generator=gen)]
    gs = [torch.randn(24, 260, 14, 14, generator=gen), torch.randn(2, 16, 56, 56, generator=gen)]
    refs = [O.factor_ema(O.conv2d_A([x], (3, 3), (1, 1), (1, 1), False), None, 0.95) for x in xs]
    refs += [O.factor_ema(O.conv2d_G([g]), None, 0.95) for g in gs]
    for name, env in (('KFB_SYRK_STREAMK', '1'), ('KFB_SYRK_STREAMK', '0'), ('KFB_SYRK_DYNAMIC', '1')):
        monkeypatch.setenv(name, env)     # the last round: CTAs claim items from a shared counter
        jobs, states = [], []
        for x in xs:
            d, segs = conv_segments(x, (3, 3), (1, 1), (1, 1), False)
            states.append(alloc2d(d, d, DEV)[1])
            jobs.append(N.FactorJob(d, segs, states[-1], True, True, None))
        for g in gs:
            n, c, h, w = g.shape
            S = h * w
            seg = N.Segment(N.KFB_SEG_NCHW, g.to(DEV), n, c, 1.0 / (S * S * n * S), h=h, w=w)
            states.append(alloc2d(c, c, DEV)[1])
            jobs.append(N.FactorJob(c, [seg], states[-1], True, True, None))
        o.factor_update(jobs, 0.95)
        torch.cuda.synchronize()
        for st, ref in zip(states, refs):
            assert_factor_close(st, ref)


def test_last_arriver_finish_repeats_and_matches_finalize(monkeypatch):
    """Split-K jobs are finished inside the SYRK by the last K chunk of each 256 x 256 block to arrive, which
    also re-zeroes the block's accumulator and counter.  The factor workspace of CudaOps is dedicated, so from
    the second update on the library clears nothing itself: four updates in a row (running averages chained,
    packed triangles exported) must each give the oracle's factor, a caller-supplied undedicated workspace
    full of garbage must too, and KFB_SYRK_LAST_ARRIVER=0 (scratch + finalize_kernel) must agree."""
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    o.set_mma_mode(1)
    gen = torch.Generator().manual_seed(77)
    # long-K single segment (several chunks), two segments, 300 = ragged 128-tiles, 64x64 map = TMA-fed
    xs = [torch.randn(10, 64, 28, 28, generator=gen)]                       # im2col 3x3: d = 576, K = 7840
    gs = [torch.randn(24, 300, 14, 14, generator=gen), torch.randn(3, 130, 64, 64, generator=gen)]
    g_two = [torch.randn(9, 96, 28, 28, generator=gen), torch.randn(5, 96, 28, 28, generator=gen)]

    def build(states, tris):
        jobs = []
        for x in xs:
            d, segs = conv_segments(x, (3, 3), (1, 1), (1, 1), True)
            jobs.append((d, segs))
        for g in gs:
            n, c, h, w = g.shape
            S = h * w
            jobs.append((c, [N.Segment(N.KFB_SEG_NCHW, g.to(DEV), n, c, 1.0 / (S * S * n * S), h=h, w=w)]))
        S = 28 * 28
        wgt = 1.0 / (S * S * 14 * S)
        jobs.append((96, [N.Segment(N.KFB_SEG_NCHW, g.to(DEV), g.shape[0], 96, wgt, h=28, w=28) for g in g_two]))
        return [N.FactorJob(d, segs, st, False, True, tri) for (d, segs), st, tri in zip(jobs, states, tris)]

    raw = [O.conv2d_A([x], (3, 3), (1, 1), (1, 1), True) for x in xs] + [O.conv2d_G([g]) for g in gs]
    raw.append(O.conv2d_G(g_two))
    dims = [r.shape[0] for r in raw]
    results = {}
    for env in ('1', '0'):
        monkeypatch.setenv('KFB_SYRK_LAST_ARRIVER', env)
        states = [alloc2d(d, d, DEV)[1] for d in dims]
        tris = [torch.full((d * (d + 1) // 2,), float('nan'), device=DEV) for d in dims]
        for st in states:
            st.copy_(torch.eye(st.shape[0]))
        refs = [torch.eye(d, dtype=torch.float64) for d in dims]
        jobs = build(states, tris)
        for step in range(4):
            o.factor_update(jobs, 0.9)
            torch.cuda.synchronize()
            refs = [O.factor_ema(r, ref.float(), 0.9).double() for r, ref in zip(raw, refs)]
            for st, tri, ref in zip(states, tris, refs):
                assert_factor_close(st, ref.float())
                assert torch.equal(tri.cpu(), triu_pack(st.cpu()))
        results[env] = [st.clone() for st in states]
    for a, b in zip(results['1'], results['0']):
        assert_factor_close(a, b, tol=1e-5)
    # write_state = False: only the packed triangle receives the new average, the state stays as it was
    monkeypatch.setenv('KFB_SYRK_LAST_ARRIVER', '1')
    g = gs[0]
    n, c, h, w = g.shape
    S = h * w
    seg = N.Segment(N.KFB_SEG_NCHW, g.to(DEV), n, c, 1.0 / (S * S * n * S), h=h, w=w)
    st = alloc2d(c, c, DEV)[1]
    st.copy_(torch.eye(c))
    tri = torch.full((c * (c + 1) // 2,), float('nan'), device=DEV)
    o.factor_update([N.FactorJob(c, [seg], st, False, False, tri)], 0.9)
    torch.cuda.synchronize()
    assert torch.equal(st.cpu(), torch.eye(c))
    want = O.factor_ema(raw[1], torch.eye(c), 0.9)
    full = torch.zeros(c, c)
    iu = torch.triu_indices(c, c)
    full[iu[0], iu[1]] = tri.cpu()
    full = full + full.triu(1).t()
    assert_factor_close(full, want)
    # an ordinary (undedicated) workspace holding garbage: cleared by the call itself
    monkeypatch.setenv('KFB_SYRK_LAST_ARRIVER', '1')
    states = [alloc2d(d, d, DEV)[1] for d in dims]
    for st in states:
        st.copy_(torch.eye(st.shape[0]))
    jobs = build(states, [None] * len(dims))
    cjobs, keep = o._factor_cjobs(jobs)
    nbytes = o.lib.kfb_factor_workspace_bytes(cjobs, len(jobs))
    ws = torch.full((nbytes,), 0x7f, dtype=torch.uint8, device=DEV)
    for step in range(2):
        o._call('kfb_factor_update', cjobs, len(jobs), 0.9, ws.data_ptr(), ws.numel(),
                torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        ws[nbytes // 2:].fill_(0x55)            # someone else scribbles on it between the calls
    refs = [torch.eye(d) for d in dims]
    for step in range(2):
        refs = [O.factor_ema(r, ref, 0.9) for r, ref in zip(raw, refs)]
    for st, ref in zip(states, refs):
        assert_factor_close(st, ref)


def test_nchw_tma_and_gather_paths_agree(monkeypatch):
    """fp32 NCHW operands with 16-byte image rows are fetched by TMA (per-image 32-row blocks, tail
    zero-filled); KFB_SYRK_TMA=0 and a capture that is not 16-byte aligned use the gathering
    producers.  All three must give the oracle's factor, across chunked and two-segment jobs."""
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    o.set_mma_mode(1)
    gen = torch.Generator().manual_seed(21)
    for n, c, h, w in [(24, 260, 14, 14), (3, 129, 6, 6), (7, 128, 28, 28)]:
        S = h * w
        g1 = torch.randn(n, c, h, w, generator=gen)
        g2 = torch.randn(n + 1, c, h, w, generator=gen)
        ref = O.factor_ema(O.conv2d_G([g1, g2]), None, 0.95)
        flat = torch.empty(g1.numel() + 1, device=DEV)          # misaligned copy: base + 4 bytes
        flat[1:].copy_(g1.reshape(-1))
        g1_off = flat[1:].view(n, c, h, w)
        assert g1_off.data_ptr() % 16 == 4
        results = []
        for env, first in (('1', g1.to(DEV)), ('0', g1.to(DEV)), ('1', g1_off)):
            monkeypatch.setenv('KFB_SYRK_TMA', env)
            wgt = 1.0 / (S * S * (2 * n + 1) * S)           # rows of both calls are averaged together
            segs = [N.Segment(N.KFB_SEG_NCHW, first, n, c, wgt, h=h, w=w),
                    N.Segment(N.KFB_SEG_NCHW, g2.to(DEV), n + 1, c, wgt, h=h, w=w)]
            buf, state = alloc2d(c, c, DEV)
            o.factor_update([N.FactorJob(c, segs, state, True, True, None)], 0.95)
            torch.cuda.synchronize()
            assert_factor_close(state, ref)
            results.append(state.clone())
        # same TF32 products, same fp32 accumulation order inside a tile: only the reduction order
        # of the K chunks (atomics) differs
        assert_factor_close(results[0], results[1], tol=1e-5)
        assert_factor_close(results[0], results[2], tol=1e-5)


@pytest.mark.parametrize('rows,feat,bias', [(32, 2048, True), (20, 13, True), (7, 100, False), (1, 5, True),
                                            (300, 65, False)])
def test_linear_factor_rows(ops, rows, feat, bias):
    a = torch.randn(rows, feat, generator=torch.Generator().manual_seed(3))
    seg = N.Segment(N.KFB_SEG_ROWS, a.to(DEV), rows, feat, 1.0 / rows, has_bias=bias, ld=feat)
    d = feat + int(bias)
    state, tri, ref = run_both(ops, d, [seg])
    assert_factor_close(state, ref)
    assert_factor_close(state, O.factor_ema(O.linear_A([a], bias), None, 0.95))


def test_linear_rows_tensor_core_and_simt_paths_agree(monkeypatch):
    """Row-major (Linear) captures reach the tcgen05 SYRK through a transposed fp32 staging copy
    (pad_kernel variant 5, bias = a row of ones); KFB_SYRK_ROWS_TC=0 keeps them on the SIMT kernel.
    Long K (split into chunks, finished by the last arriver), ragged sizes, a pitch, 16-bit rows."""
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    o.set_mma_mode(1)
    gen = torch.Generator().manual_seed(8)
    cases = [(20000, 300, True, torch.float32), (700, 651, True, torch.float32), (33, 130, False, torch.float16),
             (5, 7, True, torch.float32), (9000, 64, False, torch.bfloat16)]
    for rows, feat, bias, dtype in cases:
        wide = torch.randn(rows, feat + 3, generator=gen).to(dtype)
        x = wide.to(DEV)[:, :feat]                                       # pitch feat + 3
        d = feat + int(bias)
        ref = O.factor_ema(O.linear_A([wide[:, :feat].float()], bias), None, 0.95)
        got = {}
        for env in ('1', '0'):
            monkeypatch.setenv('KFB_SYRK_ROWS_TC', env)
            seg = N.Segment(N.KFB_SEG_ROWS, x, rows, feat, 1.0 / rows, has_bias=bias, ld=feat + 3)
            buf, state = alloc2d(d, d, DEV)
            tri = torch.full((d * (d + 1) // 2,), float('nan'), device=DEV)
            o.factor_update([N.FactorJob(d, [seg], state, True, True, tri)], 0.95)
            torch.cuda.synchronize()
            assert_factor_close(state, ref)
            assert torch.equal(tri.cpu(), triu_pack(state.cpu()))
            got[env] = state.clone()
        assert_factor_close(got['1'], got['0'])


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16], ids=['fp16', 'bf16'])
def test_half_precision_captures(ops, dtype):
    """AMP captures (fp16 / bf16) are read natively: same factor as for the widened fp32 copy."""
    g = torch.Generator().manual_seed(21)
    x = torch.randn(4, 12, 14, 14, generator=g).to(dtype)          # im2col, 14x14 map
    d, segs = conv_segments(x, (3, 3), (1, 1), (1, 1), True)
    state, _, _ = run_both(ops, d, segs)
    assert_factor_close(state, O.factor_ema(O.conv2d_A([x.float()], (3, 3), (1, 1), (1, 1), True), None, 0.95))
    go = torch.randn(3, 130, 8, 8, generator=g).to(dtype)           # NCHW, 8-byte vector path
    seg = N.Segment(N.KFB_SEG_NCHW, go.to(DEV), 3, 130, 1.0 / (64.0 * 64 * 3 * 64), h=8, w=8)
    state, _, _ = run_both(ops, 130, [seg])
    assert_factor_close(state, O.factor_ema(O.conv2d_G([go.float()]), None, 0.95))
    go7 = torch.randn(3, 40, 7, 7, generator=g).to(dtype)           # NCHW, scalar path (49 not /4)
    seg = N.Segment(N.KFB_SEG_NCHW, go7.to(DEV), 3, 40, 1.0 / (49.0 * 49 * 3 * 49), h=7, w=7)
    state, _, _ = run_both(ops, 40, [seg])
    assert_factor_close(state, O.factor_ema(O.conv2d_G([go7.float()]), None, 0.95))
    a = torch.randn(48, 65, generator=g).to(dtype)                  # Linear rows (SIMT path)
    seg = N.Segment(N.KFB_SEG_ROWS, a.to(DEV), 48, 65, 1.0 / 48, has_bias=True, ld=65)
    state, _, _ = run_both(ops, 66, [seg])
    assert_factor_close(state, O.factor_ema(O.linear_A([a.float()], True), None, 0.95))


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16], ids=['fp16', 'bf16'])
def test_16bit_captures_feed_f16_mmas(monkeypatch, dtype):
    """AMP captures that TMA can fetch (NCHW / Linear rows) stay 16-bit all the way into shared memory and are
    multiplied by kind::f16 MMAs: the products are exact, so against the float64 factor of the SAME 16-bit
    values the error is fp32 accumulation only - fifty times inside the TF32 tolerance.  In place (S % 8 == 0),
    through the 16-bit copy (14 x 14, 7 x 7 maps), split along K (last-arriver finish), two segments, rows
    with a bias, and k x k convs through 16-bit staging planes; KFB_SYRK_HALF=0 (fp32 staging copies + kind::tf32) must agree within the TF32 tolerance."""
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    o.set_mma_mode(1)
    gen = torch.Generator().manual_seed(5)
    jobs = []
    for shape in [(10, 200, 28, 28), (6, 130, 14, 14), (5, 300, 7, 7), (2, 64, 56, 56)]:
        g = torch.randn(*shape, generator=gen).to(dtype)
        n, c, h, w = shape
        S = h * w
        jobs.append((c, [(N.KFB_SEG_NCHW, g, dict(n=n, c=c, weight=1.0 / (S * S * n * S), h=h, w=w))], O.conv2d_G([g.float()])))
    g2 = [torch.randn(9, 96, 28, 28, generator=gen).to(dtype), torch.randn(5, 96, 28, 28, generator=gen).to(dtype)]
    S = 28 * 28
    jobs.append((96, [(N.KFB_SEG_NCHW, g, dict(n=g.shape[0], c=96, weight=1.0 / (S * S * 14 * S), h=28, w=28)) for g in g2],
                 O.conv2d_G([g.float() for g in g2])))
    a = torch.randn(700, 650, generator=gen).to(dtype)
    jobs.append((651, [(N.KFB_SEG_ROWS, a, dict(n=700, c=650, weight=1.0 / 700, has_bias=True, ld=650))], O.linear_A([a.float()], True)))
    # k x k convs: the per-kx staging planes keep the 16-bit type (eight pixels per 16-byte group); every
    # CONV_CASES geometry: odd widths, strides, 7x7 taps, a 7x7 map, K chunks, bias features (two 16-bit ones per word)
    for n, c, h, w, k, st, pd, bias in CONV_CASES:
        x = torch.randn(n, c, h, w, generator=gen).to(dtype)
        d, segs = conv_segments(x, k, st, pd, bias)
        sg = segs[0]
        jobs.append((d, [(N.KFB_SEG_IM2COL, x, dict(n=n, c=c, weight=sg.weight, has_bias=bias, h=h, w=w, kernel=k, stride=st,
                                                    padding=pd))],
                     O.conv2d_A([x.float()], k, st, pd, bias)))
    results = {}
    for env in ('1', '0'):
        monkeypatch.setenv('KFB_SYRK_HALF', env)
        fjobs, states = [], []
        for d, segs, _ in jobs:
            st = alloc2d(d, d, DEV)[1]
            states.append(st)
            fjobs.append(N.FactorJob(d, [N.Segment(kind, t.to(DEV), **kw) for kind, t, kw in segs], st, True, True, None))
        o.factor_update(fjobs, 0.95)
        torch.cuda.synchronize()
        for st, (d, segs, raw) in zip(states, jobs):
            assert_factor_close(st, O.factor_ema(raw, None, 0.95), tol=2e-5 if env == '1' else 1e-3)
        results[env] = [st.clone() for st in states]
    for x, y in zip(results['1'], results['0']):
        assert_factor_close(x, y)


def test_multi_segment_and_running_average(ops):
    """LinearMultiLayer: sum over calls, each with its own 1/rows; then a second EMA update."""
    gen = torch.Generator().manual_seed(11)
    calls = [torch.randn(r, 37, generator=gen) for r in (20, 20, 5, 33)]
    segs = [N.Segment(N.KFB_SEG_ROWS, c.to(DEV), c.shape[0], 37, 1.0 / c.shape[0], has_bias=True, ld=37)
            for c in calls]
    state, _, ref = run_both(ops, 38, segs)
    assert_factor_close(state, ref)
    assert_factor_close(state, O.factor_ema(O.linear_multi_A(calls, True), None, 0.95))
    # second update starts from the previous state (not the identity)
    state2, _, ref2 = run_both(ops, 38, segs[:2], alpha=0.5, old=ref)
    assert_factor_close(state2, ref2)
    expect = O.factor_ema(O.linear_multi_A(calls[:2], True), ref.clone(), 0.5)
    assert_factor_close(state2, expect)


def test_row_pitch_and_noncontiguous_rows(ops):
    base = torch.randn(50, 40, generator=torch.Generator().manual_seed(5))
    view = base[:, :29]                      # pitch 40, 29 features
    seg = N.Segment(N.KFB_SEG_ROWS, base.to(DEV)[:, :29], 50, 29, 1.0 / 50, ld=40)
    buf, st = alloc2d(29, 29, DEV)
    ops.factor_update([N.FactorJob(29, [seg], st, True, True, None)], 0.9)
    ref = O.factor_ema(O.linear_A([view.contiguous()], False), None, 0.9)
    assert_factor_close(st, ref)
    assert st.stride(0) == 32        # padded pitch, logical 29 x 29 view


def test_batched_jobs_match_golden_captures(ops):
    """All 12 factors of the golden SmallCNN step in ONE native call; compared with the
    factors the unmodified reference produced."""
    rec = load_golden('cnn_single.pt')
    import models
    import distributed_kfac_pytorch_b200 as kfac
    N.set_ops(ops)
    model = models.make_cnn().to(DEV)
    pre = kfac.KFAC(model, **rec['kfac_kwargs'])
    step = rec['steps'][0]
    for i, l in enumerate(pre.layers):
        l.a_inputs = [t.to(DEV) for t in step['a_inputs'][i]]
        l.g_outputs = [t.to(DEV) for t in step['g_outputs'][i]]
    pre.compute_factors(alpha=0.95)
    for i, l in enumerate(pre.layers):
        assert_factor_close(l.state['A'], step['A'][i])
        assert_factor_close(l.state['G'], step['G'][i])


def test_tri_pack_unpack_roundtrip(ops):
    pairs, refs = [], []
    for d in (1, 5, 32, 33, 147, 300):
        m = torch.randn(d, d, generator=torch.Generator().manual_seed(d))
        m = m + m.t()
        _, st = alloc2d(d, d, DEV)
        st.copy_(m)
        tri = torch.empty(d * (d + 1) // 2, device=DEV)
        pairs.append((tri, st))
        refs.append(m)
    ops.tri_pack(pairs)
    for (tri, st), m in zip(pairs, refs):
        torch.testing.assert_close(tri.cpu(), triu_pack(m), rtol=0, atol=0)
        st.fill_(float('nan'))
    ops.tri_unpack(pairs, 0.25)
    for (tri, st), m in zip(pairs, refs):
        torch.testing.assert_close(st.cpu(), m * 0.25, rtol=0, atol=0)


@pytest.mark.parametrize('d', [1, 7, 64, 147, 500])
def test_syevd(ops, d):
    g = torch.Generator().manual_seed(d)
    x = torch.randn(max(d // 2, 1), d, generator=g)      # rank-deficient: exercises the clip at 0
    f = x.t() @ x / x.shape[0]
    _, F_ = alloc2d(d, d, DEV)
    F_.copy_(f)
    _, E = alloc2d(d, d, DEV)
    _, Q = alloc2d(d, d, DEV)
    w = torch.empty(d, device=DEV)
    info = torch.ones(1, dtype=torch.int32, device=DEV)
    ops.syevd(F_, E, Q, w, info)
    torch.cuda.synchronize()
    assert info.item() == 0
    torch.testing.assert_close(F_.cpu(), f, rtol=0, atol=0)       # input preserved
    Qc, Ec, wc = Q.cpu().double(), E.cpu().double(), w.cpu().double()
    assert (wc >= 0).all()
    torch.testing.assert_close(Ec, Qc.t(), rtol=0, atol=0)
    torch.testing.assert_close(Qc.t() @ Qc, torch.eye(d, dtype=torch.double), rtol=0, atol=1e-4)
    recon = Qc @ torch.diag(wc) @ Qc.t()
    scale = f.abs().max().item()
    assert (recon - f.double()).abs().max().item() <= 1e-4 * max(scale, 1e-6)
    ref_w = torch.linalg.eigvalsh(f.double()).clamp_min(0)
    torch.testing.assert_close(wc, ref_w, rtol=1e-3, atol=1e-5 * max(scale, 1e-6))


def test_syevd_batch_matches_single(ops):
    """The multi-lane batch gives the same decompositions as one-at-a-time calls."""
    dims = [3, 64, 64, 147, 256, 33, 512, 64, 100, 8, 300, 1]
    single, batch, factors = [], [], []
    for i, d in enumerate(dims):
        x = torch.randn(d + 3, d, generator=torch.Generator().manual_seed(50 + i))
        f = (x.t() @ x / x.shape[0]).to(DEV)
        _, F_ = alloc2d(d, d, DEV)
        F_.copy_(f)
        factors.append(F_)
        for lst in (single, batch):
            lst.append((F_, alloc2d(d, d, DEV)[1], alloc2d(d, d, DEV)[1], torch.empty(d, device=DEV),
                        torch.ones(1, dtype=torch.int32, device=DEV)))
    for job in single:
        ops.syevd(*job)
    ops.syevd_batch(batch)
    torch.cuda.synchronize()
    for (f, E1, Q1, w1, i1), (_, E2, Q2, w2, i2) in zip(single, batch):
        assert i1.item() == 0 and i2.item() == 0
        torch.testing.assert_close(w2, w1, rtol=1e-5, atol=1e-6 * w1.abs().max().item())
        torch.testing.assert_close(E2, Q2.t(), rtol=0, atol=0)
        d = f.shape[0]
        recon = (Q2.double() * w2.double()) @ Q2.double().t()
        assert (recon - f.double()).abs().max().item() <= 2e-4 * max(f.abs().max().item(), 1e-6)


@pytest.mark.parametrize('d', [1, 5, 64, 147, 400])
def test_cholesky_inverse(ops, d):
    x = torch.randn(2 * d + 3, d, generator=torch.Generator().manual_seed(d))
    f = x.t() @ x / x.shape[0]
    _, F_ = alloc2d(d, d, DEV)
    F_.copy_(f)
    buf, out = alloc2d(d, d, DEV)
    buf.fill_(float('nan'))
    info = torch.ones(1, dtype=torch.int32, device=DEV)
    ops.cholesky_inverse(F_, 0.003, out, info)
    torch.cuda.synchronize()
    assert info.item() == 0
    ref = O.damped_inverse(f, 0.003)
    torch.testing.assert_close(out.cpu(), ref, rtol=2e-3, atol=2e-4 * ref.abs().max().item())
    assert torch.equal(out, out.t())
    torch.testing.assert_close(F_.cpu(), f, rtol=0, atol=0)


def test_transpose_and_outer_recip(ops):
    a = torch.randn(70, 45)
    _, src = alloc2d(70, 45, DEV)
    src.copy_(a)
    _, dst = alloc2d(45, 70, DEV)
    ops.transpose(src, dst)
    torch.testing.assert_close(dst.cpu(), a.t(), rtol=0, atol=0)
    dG, dA = torch.rand(33), torch.rand(70)
    _, out = alloc2d(33, 70, DEV)
    ops.outer_recip(dG.to(DEV), dA.to(DEV), 0.003, out)
    torch.testing.assert_close(out.cpu(), O.outer_recip(dG, dA, 0.003), rtol=1e-6, atol=0)


GEMM_SHAPES = [(128, 128, 32), (128, 128, 64), (256, 128, 512), (64, 64, 64), (147, 64, 147), (1000, 2049, 32),
               (300, 130, 577), (1, 1, 1), (129, 257, 33), (512, 4608, 512), (2048, 512, 2048)]


@pytest.mark.parametrize('shape', GEMM_SHAPES, ids=lambda s: 'x'.join(map(str, s)))
def test_gemm_tn(ops, shape):
    """D = A B^T against float64; TF32 keeps 10 mantissa bits per operand, so the bound is
    relative to |A||B| row norms (about 1e-3), well inside the preconditioned-gradient tolerance."""
    M, Nn, K = shape
    g = torch.Generator().manual_seed(M * 7 + Nn * 3 + K)
    A = torch.randn(M, K, generator=g)
    B = torch.randn(Nn, K, generator=g)
    mul = torch.rand(M, Nn, generator=g) + 0.5
    _, a = alloc2d(M, K, DEV, zero=True); a.copy_(A)
    _, b = alloc2d(Nn, K, DEV, zero=True); b.copy_(B)
    _, m = alloc2d(M, Nn, DEV, zero=True); m.copy_(mul)
    buf, d = alloc2d(M, Nn, DEV)
    buf.fill_(float('nan'))
    ops.gemm_tn(a, b, d, m)
    torch.cuda.synchronize()
    ref = (A.double() @ B.double().t()) * mul.double()
    bound = 2e-3 * torch.outer(A.double().norm(dim=1), B.double().norm(dim=1)) * mul.double() + 1e-6
    err = (d.cpu().double() - ref).abs()
    assert torch.isfinite(d).all()
    worst = (err / bound).max().item()
    assert worst <= 1.0, 'gemm error %.3g x bound' % worst
    if ops.get_mma_mode() == 0:
        torch.testing.assert_close(d.cpu(), ref.float(), rtol=1e-4, atol=1e-5 * K ** 0.5 * 4)
    if buf.shape[1] > Nn:     # the pitch padding must not be written
        assert torch.isnan(buf[:, Nn:]).all()


@pytest.mark.parametrize('shape', GEMM_SHAPES, ids=lambda s: 'x'.join(map(str, s)))
def test_gemm_tn_3xtf32(shape):
    """mma mode 2: every product is A_lo B_hi + A_hi B_lo + A_hi B_hi (the low parts are made in shared
    memory by the kernel's split warps).  Same tiles, same TMA traffic, three times the MMAs; the error
    bound against float64 is 400 times tighter than single-pass TF32's."""
    M, Nn, K = shape
    g = torch.Generator().manual_seed(M * 7 + Nn * 3 + K)
    A = torch.randn(M, K, generator=g)
    B = torch.randn(Nn, K, generator=g)
    mul = torch.rand(M, Nn, generator=g) + 0.5
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    _, a = alloc2d(M, K, DEV, zero=True); a.copy_(A)
    _, b = alloc2d(Nn, K, DEV, zero=True); b.copy_(B)
    _, m = alloc2d(M, Nn, DEV, zero=True); m.copy_(mul)
    buf, d = alloc2d(M, Nn, DEV)
    buf.fill_(float('nan'))
    o.set_mma_mode(2)
    try:
        o.gemm_tn(a, b, d, m)
        torch.cuda.synchronize()
    finally:
        o.set_mma_mode(1)
    ref = (A.double() @ B.double().t()) * mul.double()
    bound = 5e-6 * torch.outer(A.double().norm(dim=1), B.double().norm(dim=1)) * mul.double() + 1e-7
    err = (d.cpu().double() - ref).abs()
    assert torch.isfinite(d).all()
    worst = (err / bound).max().item()
    assert worst <= 1.0, 'gemm error %.3g x bound' % worst
    if buf.shape[1] > Nn:
        assert torch.isnan(buf[:, Nn:]).all()


def make_precond_job(da, dg, bias, wshape, seed):
    g = torch.Generator().manual_seed(seed)
    A = torch.randn(2 * da, da, generator=g)
    A = A.t() @ A / (2 * da)
    G = torch.randn(2 * dg, dg, generator=g)
    G = G.t() @ G / (2 * dg)
    QA, wA = O.eigh_clipped(A)
    QG, wG = O.eigh_clipped(G)
    dGdA = O.outer_recip(wG, wA, 0.003)
    wgrad = torch.randn(*wshape, generator=g)
    bgrad = torch.randn(dg, generator=g) if bias else None
    bufs = {}

    def put(name, t):
        _, v = alloc2d(t.shape[0], t.shape[1], DEV, zero=True)
        v.copy_(t)
        bufs[name] = v
    put('QA', QA); put('EA', QA.t().contiguous()); put('QG', QG); put('EG', QG.t().contiguous())
    put('dGdA', dGdA)
    for name, shape in (('gmat', (dg, da)), ('V', (dg, da)), ('out', (dg, da)), ('T', (da, dg))):
        put(name, torch.zeros(shape))
    job = N.PrecondJob(da=da, dg=dg, has_bias=bias, **bufs)
    job.weight_grad = wgrad.to(DEV)
    job.bias_grad = bgrad.to(DEV) if bias else None
    ref = O.precondition_eigen(O.grad_matrix(wgrad, bgrad), QA, QG, dGdA=dGdA)
    return job, ref, wgrad, bgrad


PRECOND_CASES = [(27, 16, False, (16, 3, 3, 3)), (65, 10, True, (10, 64)), (147, 64, False, (64, 3, 7, 7)),
                 (256, 128, False, (128, 256, 1, 1)), (577, 130, True, (130, 64, 3, 3)), (1, 1, False, (1, 1))]


def test_precondition_chain_and_kl_clip(ops):
    jobs, refs, raw = [], [], []
    for i, (da, dg, bias, wshape) in enumerate(PRECOND_CASES):
        job, ref, wg, bg = make_precond_job(da, dg, bias, wshape, 100 + i)
        jobs.append(job); refs.append(ref); raw.append((wg, bg))
    vg = torch.zeros(1, device=DEV)
    ops.precondition(jobs, vg)
    torch.cuda.synchronize()
    vg_ref = 0.0
    for job, ref, (wg, bg) in zip(jobs, refs, raw):
        scale = ref.abs().max().item()
        # north_star: rtol 1e-2 / atol 1e-4 at the reference's gradient scale (O(1e-3) after
        # KL clipping).  These synthetic problems have |out| up to ~30, so the absolute term is
        # scaled to the problem: fp32 SIMT 1e-4 * scale, single-pass TF32 2e-3 * scale (10-bit
        # mantissas through four chained GEMMs with 1/damping = 333 amplification in between).
        atol = (1e-4 if ops.get_mma_mode() == 0 else 2e-3) * max(1.0, scale)
        torch.testing.assert_close(job.out.cpu(), ref, rtol=1e-2, atol=atol)
        vg_ref += (ref * O.grad_matrix(wg, bg)).sum().item()
    assert abs(vg.item() - vg_ref) <= 1e-2 * abs(vg_ref)
    # dot_accumulate reproduces the fused reduction
    vg2 = torch.zeros(1, device=DEV)
    items = [(j.weight_grad, j.bias_grad, j.out) for j in jobs]
    ops.dot_accumulate(items, vg2)
    assert abs(vg2.item() - vg.item()) <= 1e-4 * abs(vg.item())
    # KL-clipped write-back, no host sync inside
    lr, kl = 0.1, 0.001
    ops.scale_writeback(items, vg, lr, kl)
    nu = min(1.0, math.sqrt(kl / abs(vg.item() * lr * lr)))
    for job, ref in zip(jobs, refs):
        n_w = job.da - int(job.has_bias)
        got = job.weight_grad.cpu().reshape(job.dg, n_w)
        torch.testing.assert_close(got, nu * job.out.cpu()[:, :n_w], rtol=1e-6, atol=0)
        if job.has_bias:
            torch.testing.assert_close(job.bias_grad.cpu(), nu * job.out.cpu()[:, n_w], rtol=1e-6, atol=0)
    # vg == 0 -> no scaling (reference returns None, preconditioner.py:680-681)
    vg.zero_()
    before = jobs[0].out.clone()
    ops.scale_writeback(items[:1], vg, lr, kl)
    torch.testing.assert_close(jobs[0].weight_grad.reshape(before.shape), before, rtol=0, atol=0)


def test_precondition_chain_3xtf32_matches_fp32():
    """mma mode 2 on the whole chain (larger layers added so that every stage runs on the tensor cores): the
    result must meet the bound the exact fp32 SIMT path is held to (1e-4 * scale), twenty times tighter than
    single-pass TF32's, and agree with the SIMT path itself."""
    cases = PRECOND_CASES + [(1152, 256, False, (256, 128, 3, 3)), (513, 1024, True, (1024, 512))]
    N.set_ops(None)
    o = N.get_ops()
    o._enter(torch.zeros(1, device=DEV))
    outs = {}
    for mode in (2, 0):
        jobs, refs = [], []
        for i, (da, dg, bias, wshape) in enumerate(cases):
            job, ref, wg, bg = make_precond_job(da, dg, bias, wshape, 100 + i)
            jobs.append(job); refs.append(ref)
        vg = torch.zeros(1, device=DEV)
        o.set_mma_mode(mode)
        try:
            o.precondition(jobs, vg)
            torch.cuda.synchronize()
        finally:
            o.set_mma_mode(1)
        for job, ref in zip(jobs, refs):
            scale = max(1.0, ref.abs().max().item())
            torch.testing.assert_close(job.out.cpu(), ref, rtol=1e-2, atol=1e-4 * scale)
        outs[mode] = [j.out.cpu().clone() for j in jobs], vg.item()
    for a, b, ref in zip(outs[2][0], outs[0][0], refs):
        diff, scale = (a - b).abs().max().item(), max(1.0, ref.abs().max().item())
        assert diff <= 1e-4 * scale, 'modes 2 and 0 differ by %.3g on a scale of %.3g' % (diff, scale)
        print('3xTF32 vs fp32: %.3g of scale %.3g; fp32 vs float64 oracle: %.3g' % (diff, scale, (b - ref).abs().max().item()))
    assert abs(outs[2][1] - outs[0][1]) <= 1e-4 * abs(outs[0][1])


def test_capture_filter_and_16bit_round_trip(ops):
    """kfb_capture_filter (finite test, 1/s^2, renormalisation, skip) and kfb_narrow / kfb_widen."""
    g = torch.Generator(device=DEV).manual_seed(5)
    caps = [torch.randn(4, 6, 5, 5, device=DEV, generator=g) for _ in range(3)]
    caps += [torch.randn(64, 33, device=DEV, generator=g).half(), torch.randn(1000, device=DEV, generator=g).bfloat16()]
    caps[1].view(-1)[-1] = float('nan')                 # in the scalar tail of the scan
    caps[3].view(-1)[17] = float('inf')
    scale_dev = torch.full((1,), 8.0, device=DEV)
    scal = torch.empty(5 + 3, device=DEV)
    f1 = N.CaptureFilter(caps[:3], [100.0, 100.0, 50.0], [4.0, scale_dev, scale_dev], True, scal[0:3], scal[5:6])
    f2 = N.CaptureFilter(caps[3:4], [64.0], [2.0], True, scal[3:4], scal[6:7])
    f3 = N.CaptureFilter(caps[4:5], [10.0], [2.0], False, scal[4:5], scal[7:8])
    ops.capture_filter([f1, f2, f3])
    got = scal.cpu().tolist()
    assert got[:5] == [1 / 16.0, 0.0, 1 / 64.0, 0.0, 0.25]
    assert got[5:] == [pytest.approx(250.0 / 150.0, rel=1e-6), -1.0, 1.0]
    # a dropped segment contributes nothing even though its sums are nan; a skipped job keeps its state
    x = caps[0][:, :, 0, 0].contiguous()          # 4 x 6 rows
    bad = x.clone()
    bad[0, 0] = float('nan')
    keep, drop, skip = scal[0:1], scal[1:2], scal[6:7]
    seg_ok = N.Segment(N.KFB_SEG_ROWS, x, 4, 6, 0.25, ld=6)
    seg_bad = N.Segment(N.KFB_SEG_ROWS, bad, 4, 6, 0.25, ld=6)
    seg_ok.dev_scale, seg_bad.dev_scale = keep, drop
    _, st = alloc2d(6, 6, DEV, zero=True)
    ops.factor_update([N.FactorJob(6, [seg_ok, seg_bad], st, False, True, None)], 0.5)
    ref = 0.5 * (0.25 / 16.0) * (x.t() @ x)
    assert_factor_close(st, ref)      # row-major captures run in TF32 on the tensor-core path
    st2 = torch.full((6, 8), 3.0, device=DEV)[:, :6]
    job = N.FactorJob(6, [seg_ok], st2, False, True, None)
    job.dev_rescale = skip
    ops.factor_update([job], 0.5)
    assert torch.equal(st2, torch.full_like(st2, 3.0))
    for dt in (torch.float16, torch.bfloat16):
        src = torch.randn(1003, device=DEV, generator=g) * 3
        n16 = torch.empty(1003, device=DEV, dtype=dt)
        ops.narrow(src, n16)
        assert torch.equal(n16, src.to(dt))
        back = torch.empty(1003, device=DEV)
        ops.widen(n16, back)
        assert torch.equal(back, src.to(dt).float())


def test_layer_utils_api(ops):
    """kfac.layers.utils: the stand-alone helpers of the reference (kfac/layers/utils.py), on the kernels."""
    from distributed_kfac_pytorch_b200.layers import utils as U
    g = torch.Generator(device=DEV).manual_seed(9)
    a = torch.randn(50, 13, device=DEV, generator=g)
    b = torch.randn(50, 13, device=DEV, generator=g)
    torch.testing.assert_close(U.get_cov(a), O.second_moment(a.cpu()).to(DEV), rtol=2e-3, atol=1e-4)
    torch.testing.assert_close(U.get_cov(a, b, scale=7.0), (a.t() @ (b / 7.0)), rtol=2e-3, atol=2e-3)
    f = U.get_cov(a) + torch.eye(13, device=DEV)
    q, d = U.get_eigendecomp(f, concat=False)
    torch.testing.assert_close((q * d) @ q.t(), f, rtol=1e-4, atol=1e-4)
    assert U.get_eigendecomp(f).shape == (13, 14)
    torch.testing.assert_close(U.get_inverse(f, damping=0.1) @ (f + 0.1 * torch.eye(13, device=DEV)),
                               torch.eye(13, device=DEV), rtol=1e-3, atol=1e-3)
    tri = U.get_triu(f)
    assert tri.numel() == 13 * 14 // 2
    torch.testing.assert_close(U.fill_triu([13, 13], tri), f, rtol=0, atol=0)
    assert U.append_bias_ones(a).shape == (50, 14) and bool((U.append_bias_ones(a)[:, -1] == 1).all())
    cur = torch.eye(13, device=DEV)
    U.update_running_avg(f, cur, 0.9)
    torch.testing.assert_close(cur, 0.9 * torch.eye(13, device=DEV) + 0.1 * f, rtol=1e-5, atol=1e-6)
    with pytest.raises(ValueError):
        U.get_cov(a[0])


def test_native_errors_are_loud(ops):
    _, st = alloc2d(8, 8, DEV)
    bad = N.Segment(N.KFB_SEG_ROWS, torch.zeros(4, 5, device=DEV), 4, 5, 1.0, ld=5)   # 5 features, d = 8
    with pytest.raises(N.NativeError, match='feature count'):
        ops.factor_update([N.FactorJob(8, [bad], st, True, True, None)], 0.9)
    with pytest.raises(N.NativeError):
        ops.transpose(torch.zeros(3, 3), torch.zeros(3, 3, device=DEV))
