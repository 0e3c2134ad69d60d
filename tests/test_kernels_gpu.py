"""
GPU parity tests, one per C-ABI entry point: every CUDA kernel of libpctts.so (called through
percivaltts_b200.ops.CudaOps, i.e. through the C ABI) against the plain-torch CPU definition of the same operator
(tests/cpu_ops_double.py, computed in float64) on seeded inputs.  Tolerances are fp32 round-off levels
(north_star: 1e-3 relative for the step; kernels are held to 1e-5..1e-4).
"""
import numpy as np
import pytest
import torch

from cpu_ops_double import CpuOpsDouble
from percivaltts_b200 import ops as OPS

pytestmark = pytest.mark.gpu

REF = CpuOpsDouble(torch.float64)


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed + 17 * len(shape) + sum(shape))
    return (torch.randn(*shape, generator=g, dtype=torch.float64) * scale)


def dev(t):
    return t.to(torch.float32).cuda() if t.is_floating_point() else t.cuda()


def close(got, want, tol=2e-5, what=''):
    got = got.detach().cpu().double()
    scale = float(want.abs().max()) + 1e-30
    err = float((got - want).abs().max()) / scale
    assert err < tol, (what, err)


GEMM_CASES = [
    # M, N, K, transA, transB
    (300, 70, 45, False, False), (129, 64, 16, False, True), (64, 33, 257, True, False), (17, 5, 9, True, True),
    (1000, 256, 300, False, False), (256, 1, 515, True, False), (515, 256, 1, False, True), (1, 1, 40, False, False),
    # Dense(1) heads: matrix-vector, outer product, weighted column sum (split over k) fast paths of gemm_simt.cu
    (1003, 1, 256, False, False), (1003, 1, 512, False, True), (777, 256, 1, False, False), (300, 33, 1, True, True),
    (256, 1, 3001, True, False), (512, 1, 4000, True, True),
]


@pytest.mark.parametrize('M,N,K,ta,tb', GEMM_CASES)
def test_gemm_plain_and_epilogues(cuda_ops_simt, M, N, K, ta, tb):
    A = rnd(K, M) if ta else rnd(M, K)
    B = rnd(N, K, seed=1) if tb else rnd(K, N, seed=1)
    bias, aux, addmod = rnd(N, seed=2), rnd(M, N, seed=3), rnd(max(M // 3, 1), N, seed=4)
    for flags, kw in ((0, {}), (OPS.EPI_LRELU, dict(bias=bias)), (OPS.EPI_SIGMOID, dict(bias=bias)),
                      (0, dict(aux=aux)), (OPS.EPI_LRELU, dict(addmod=addmod)), (OPS.EPI_ACCUM, dict(bias=bias))):
        C0 = rnd(M, N, seed=5)
        want = C0.clone()
        REF.gemm(A, B, want, transA=ta, transB=tb, alpha=0.7, flags=flags, **kw)
        got = dev(C0)
        cuda_ops_simt.gemm(dev(A), dev(B), got, transA=ta, transB=tb, alpha=0.7, flags=flags,
                      **dict((k, dev(v)) for k, v in kw.items()))
        close(got, want, 3e-5, ('gemm', flags, list(kw)))


def test_gemm_segmented_views_conv1d_im2col(cuda_ops_simt):
    """Conv1D 'same' as a GEMM over the virtual im2col view of a zero-padded buffer, forward and weight gradient."""
    Bn, T, C, W, H = 3, 29, 11, 5, 13
    xp = torch.zeros(Bn, T + W - 1, C, dtype=torch.float64)
    xp[:, 2:2 + T] = rnd(Bn, T, C)
    k = rnd(W * C, H, seed=1)
    view = lambda t: t.as_strided((Bn, T, W * C), ((T + W - 1) * C, C, 1))
    want = torch.zeros(Bn * T, H, dtype=torch.float64)
    REF.gemm(view(xp), k, want)
    got = torch.zeros(Bn * T, H, device='cuda')
    xpd = dev(xp)
    cuda_ops_simt.gemm(view(xpd), dev(k), got)
    close(got, want, 2e-5, 'conv1d fwd')
    dz = rnd(Bn * T, H, seed=2)
    wantw = torch.zeros(W * C, H, dtype=torch.float64)
    REF.gemm(view(xp), dz, wantw, transA=True)
    gotw = torch.zeros(W * C, H, device='cuda')
    cuda_ops_simt.gemm(view(xpd), dev(dz), gotw, transA=True)
    close(gotw, wantw, 2e-5, 'conv1d wgrad')
    # strided C / sliced 3-D operands (LSTM recurrent-kernel gradient pattern)
    h = rnd(Bn, T, 2 * H, seed=3)
    dzl = rnd(Bn, T, 4 * H, seed=4)
    wantu = torch.zeros(H, 4 * H, dtype=torch.float64)
    REF.gemm(h[:, 0:T - 1, H:], dzl[:, 1:T], wantu, transA=True)
    gotu = torch.zeros(H, 4 * H, device='cuda')
    cuda_ops_simt.gemm(dev(h)[:, 0:T - 1, H:], dev(dzl)[:, 1:T], gotu, transA=True)
    close(gotu, wantu, 2e-5, 'lstm dU')
    out = rnd(Bn * T, 7, seed=5)
    wanto = out.clone()
    REF.gemm(h.view(Bn * T, 2 * H), rnd(2 * H, 1, seed=6), wanto[:, 2:3], bias=rnd(1, seed=7))
    goto = dev(out)
    cuda_ops_simt.gemm(dev(h).view(Bn * T, 2 * H), dev(rnd(2 * H, 1, seed=6)), goto[:, 2:3], bias=dev(rnd(1, seed=7)))
    close(goto, wanto, 2e-5, 'strided C')


CONV_CASES = [(1, 4, 5, 5), (4, 4, 5, 5), (4, 1, 5, 5), (1, 2, 3, 3), (2, 2, 3, 3), (2, 1, 3, 3), (3, 5, 3, 5)]


@pytest.mark.parametrize('CIN,COUT,KH,KW', CONV_CASES)
@pytest.mark.parametrize('N,T,F', [(2, 37, 129), (1, 5, 17), (3, 16, 65), (4, 100, 130)])
def test_conv2d_fwd_flip_wgrad(cuda_ops, CIN, COUT, KH, KW, N, T, F):
    x, w, b = rnd(N, T, F, CIN), rnd(KH, KW, CIN, COUT, seed=1), rnd(COUT, seed=2)
    aux = rnd(N, T, F, COUT, seed=3)
    for flags, kw in ((0, {}), (OPS.EPI_LRELU, dict(bias=b)), (0, dict(aux=aux)), (OPS.EPI_ACCUM, {})):
        y0 = rnd(N, T, F, COUT, seed=4)
        want = y0.clone()
        REF.conv2d(x, w, want, flags=flags, **kw)
        got = dev(y0)
        cuda_ops.conv2d(dev(x), dev(w), got, flags=flags, **dict((k, dev(v)) for k, v in kw.items()))
        close(got, want, 2e-5, ('conv2d', flags, list(kw)))
    # in-place tangent form: aux aliases y
    want = aux.clone()
    REF.conv2d(x, w, want, aux=want)
    got = dev(aux)
    cuda_ops.conv2d(dev(x), dev(w), got, aux=got)
    close(got, want, 2e-5, 'conv2d in-place aux')
    wt_want = torch.zeros(KH, KW, COUT, CIN, dtype=torch.float64)
    REF.conv2d_flip(w, wt_want)
    wt = torch.zeros(KH, KW, COUT, CIN, device='cuda')
    cuda_ops.conv2d_flip(dev(w), wt)
    close(wt, wt_want, 1e-7, 'flip')
    dy = rnd(N, T, F, COUT, seed=5)
    dw_want = torch.zeros(KH, KW, CIN, COUT, dtype=torch.float64)
    REF.conv2d_wgrad(x, dy, dw_want, scale=0.5)
    dw = torch.zeros(KH, KW, CIN, COUT, device='cuda')
    cuda_ops.conv2d_wgrad(dev(x), dev(dy), dw, scale=0.5)
    close(dw, dw_want, 3e-5, 'wgrad')
    # fused bias gradient over the first samples only (critic: the GP rows carry no bias gradient), accumulate form
    db_want, dw_acc = torch.zeros(COUT, dtype=torch.float64), rnd(KH, KW, CIN, COUT, seed=6)
    dw_want2 = dw_acc.clone()
    nb = max(N - 1, 1)
    REF.conv2d_wgrad(x, dy, dw_want2, accumulate=True, dbias=db_want, bias_samples=nb)
    dw2, db = dev(dw_acc), torch.zeros(COUT, device='cuda')
    cuda_ops.conv2d_wgrad(dev(x), dev(dy), dw2, accumulate=True, dbias=db, bias_samples=nb)
    close(dw2, dw_want2, 3e-5, 'wgrad accumulate')
    close(db, db_want, 3e-5, 'fused bias gradient')


@pytest.mark.parametrize('N,T,F,L', [(2, 37, 129, 8), (1, 5, 17, 3), (3, 16, 65, 1), (5, 100, 130, 9), (2, 33, 9, 4)])
def test_conv2d_wgrad_multi(cuda_ops, N, T, F, L):
    """Weight + bias gradients of a whole 5x5 stack in one launch (pctts_conv2d_wgrad_multi): first layer 1 -> 4 channels read from
    a strided feature view, the others 4 -> 4; write form with the bias over the first samples, then the accumulate form."""
    feats = rnd(N, T, F + 3, seed=1)
    xs = [feats[:, :, 1:1 + F].unsqueeze(-1)] + [rnd(N, T, F, 4, seed=10 + i) for i in range(1, L)]
    dys = [rnd(N, T, F, 4, seed=30 + i) for i in range(L)]
    nb = max(N - 1, 1)
    fd = dev(feats)
    xd = [fd[:, :, 1:1 + F].unsqueeze(-1)] + [dev(x) for x in xs[1:]]
    dyd = [dev(d) for d in dys]
    dws = [torch.full((5, 5, x.shape[3], 4), 7.0, device='cuda') for x in xs]
    dbs = [torch.full((4,), 7.0, device='cuda') for _ in xs]
    layers = [dict(x=xd[i], dy=dyd[i], dw=dws[i], dbias=dbs[i]) for i in range(L)]
    assert cuda_ops.conv2d_wgrad_multi_ok(layers)
    cuda_ops.conv2d_wgrad_multi(layers, bias_samples=nb)
    want = []
    for i in range(L):
        dw_want, db_want = torch.zeros(5, 5, xs[i].shape[3], 4, dtype=torch.float64), torch.zeros(4, dtype=torch.float64)
        REF.conv2d_wgrad(xs[i], dys[i], dw_want, dbias=db_want, bias_samples=nb)
        close(dws[i], dw_want, 3e-5, ('multi wgrad', i))
        close(dbs[i], db_want, 3e-5, ('multi bias', i))
        want.append(dw_want)
    # accumulate form without biases (the gradient-penalty rows of the critic step)
    before = [b.clone() for b in dbs]
    cuda_ops.conv2d_wgrad_multi([dict(x=xd[i], dy=dyd[i], dw=dws[i]) for i in range(L)], scale=0.5, accumulate=True)
    for i in range(L):
        close(dws[i], 1.5 * want[i], 3e-5, ('multi wgrad accumulate', i))
        assert torch.equal(dbs[i], before[i])
    # run-to-run determinism
    again = [torch.zeros_like(d) for d in dws]
    cuda_ops.conv2d_wgrad_multi([dict(x=xd[i], dy=dyd[i], dw=again[i]) for i in range(L)])
    again2 = [torch.zeros_like(d) for d in dws]
    cuda_ops.conv2d_wgrad_multi([dict(x=xd[i], dy=dyd[i], dw=again2[i]) for i in range(L)])
    assert all(torch.equal(a, b) for a, b in zip(again, again2))


@pytest.mark.parametrize('rows,C', [(4000, 256), (517, 64), (33, 8)])
def test_batchnorm_apply_kernels_write_operand_planes(rows, C):
    """bn_apply / bn_bwd_apply(planes=True) write y / dx and, in the same launch, the bf16 operand planes the consuming GEMM would
    otherwise get from pctts_split_bf16: same fp32 output as the plain kernels and bit-identical planes, and the GEMM that follows
    inside the step scope launches no split."""
    ops = OPS.CudaOps(precision='fp32')
    x, dy = dev(rnd(rows, C) * 2 + 0.3), dev(rnd(rows, C, seed=1))
    gamma, beta = dev(rnd(C, seed=2) + 1.5), dev(rnd(C, seed=3) * 0.1)
    mean, istd = dev(rnd(C, seed=4) * 0.1), dev(rnd(C, seed=5).abs() + 0.5)
    y0, y1 = torch.zeros(rows, C, device='cuda'), torch.zeros(rows, C, device='cuda')
    ops.bn_apply(x, y0, mean, istd, gamma, beta, 1e-3, 0, 1)
    w = dev(rnd(C, 16, seed=6))
    out0, out1 = torch.zeros(rows, 16, device='cuda'), torch.zeros(rows, 16, device='cuda')
    with ops.step_scope():
        ops.bn_apply(x, y1, mean, istd, gamma, beta, 1e-3, 0, 1, planes=True)
        assert torch.equal(y0, y1)
        got = ops.shadow(y1)                       # fresh: no launch
        n0 = ops.launches
        assert ops.shadow(y1) is got and ops.launches == n0
        want = OPS.Shadow(ops.device, rows, C, 3)
        ops._split(y1, want)
        assert torch.equal(got.buf, want.buf)
        n0 = ops.launches
        ops.gemm(y1, w, out1)
        n_gemm = ops.launches - n0
    ops.gemm(y0, w, out0)
    assert torch.equal(out0, out1)
    # backward apply
    ws = ops.bn_bwd_reduce(x, dy, mean, istd, gamma, beta, 1)
    dx0, dx1 = torch.zeros(rows, C, device='cuda'), torch.zeros(rows, C, device='cuda')
    ops.bn_bwd_apply(ws, x, dy, dx0, mean, istd, gamma, beta, rows, 1, None, None)
    with ops.step_scope():
        ops.bn_bwd_apply(ws, x, dy, dx1, mean, istd, gamma, beta, rows, 1, None, None, planes=True)
        assert torch.equal(dx0, dx1)
        got = ops.shadow(dx1)
        want = OPS.Shadow(ops.device, rows, C, 3)
        ops._split(dx1, want)
        assert torch.equal(got.buf, want.buf)
        # a later writer invalidates the planes again
        ops.bn_apply(x, dx1, mean, istd, gamma, beta, 1e-3, 0, 1)
        n0 = ops.launches
        ops.shadow(dx1)
        assert ops.launches == n0 + 1


def test_split_bf16_multi_matches_single_splits():
    """The weight shadows of a network after Adam are refreshed by ONE launch (pctts_split_bf16_multi): bit-identical planes."""
    import ctypes
    ops = OPS.CudaOps(precision='fp32')
    shapes = [(516, 256), (256, 256), (21 * 425, 256), (256, 129), (256, 1024), (33, 7), (8, 8)]
    mats = [dev(rnd(r, c, seed=i)) for i, (r, c) in enumerate(shapes)]
    want = [OPS.Shadow(ops.device, r, c, 3) for r, c in shapes]
    got = [OPS.Shadow(ops.device, r, c, 3) for r, c in shapes]
    for m, sh in zip(mats, want):
        sh.buf.zero_()
        ops._split(m, sh)
    for sh in got:
        sh.buf.zero_()
    jobs = (OPS.SplitJobStruct * len(mats))(*[OPS.SplitJobStruct(m.data_ptr(), sh.ptr, m.stride(0), m.shape[0], sh.ld, sh.part_stride, m.shape[1])
                                             for m, sh in zip(mats, got)])
    OPS._lib.check(ops.lib.pctts_split_bf16_multi(ctypes.cast(jobs, ctypes.c_void_p), len(mats), 3, ops._stream()), 'multi')
    for a, b in zip(got, want):
        assert torch.equal(a.buf, b.buf)


def test_grad_sync_adam_single_rank_equals_adam_kernel(cuda_ops):
    """pctts_grad_sync_adam with world = 1 (no exchange): the fused launch's optimiser half against pctts_adam, the step
    counter and the control block's sequence number advance per call (the multi-rank half runs in tests/test_dp_gpu.py)."""
    import ctypes
    n = 100003
    npad = (n + 3) // 4 * 4
    g = torch.Generator(device='cuda').manual_seed(3)
    grad = torch.zeros(npad, device='cuda')
    p = torch.randn(n, device='cuda', generator=g)
    m, v = torch.zeros(n, device='cuda'), torch.zeros(n, device='cuda')
    step = torch.zeros(1, dtype=torch.int32, device='cuda')
    p2, m2, v2, step2 = p.clone(), m.clone(), v.clone(), step.clone()
    ctl = torch.zeros(int(cuda_ops.lib.pctts_grad_sync_ctl_bytes()) // 8, dtype=torch.int64, device='cuda')
    err = torch.zeros(1, dtype=torch.int32, device='cuda')
    gb, cb = (ctypes.c_ulonglong * 1)(grad.data_ptr()), (ctypes.c_ulonglong * 1)(ctl.data_ptr())
    for it in range(3):
        grad[:n].copy_(torch.randn(n, device='cuda', generator=g) * 0.1)
        cuda_ops.adam(p2, grad[:n], m2, v2, 1e-3, 0.5, 0.9, 1e-7, step2)
        OPS._lib.check(cuda_ops.lib.pctts_grad_sync_adam(ctypes.cast(gb, ctypes.c_void_p), ctypes.cast(cb, ctypes.c_void_p), 0, 1,
                                                         p.data_ptr(), m.data_ptr(), v.data_ptr(), n, npad, 1e-3, 0.5, 0.9, 1e-7,
                                                         step.data_ptr(), None, 0, err.data_ptr(), cuda_ops._stream()), 'grad_sync_adam')
    torch.cuda.synchronize()
    for a_, b_ in ((p, p2), (m, m2), (v, v2)):       # same expressions; the compilers' FMA contraction may differ by an ulp
        assert float((a_ - b_).abs().max()) <= 2e-7 * float(b_.abs().max())
    assert int(step) == int(step2) == 3 and int(ctl[0]) == 3 and int(err) == 0


@pytest.mark.parametrize('N,T,F', [(2, 37, 129), (1, 5, 17), (3, 40, 65)])
def test_conv2d_dgrad_reads_forward_kernel(N, T, F):
    """conv2d_dgrad (tensor-core kernel with PCTTS_EPI_WFLIP: the flipped kernel is formed while the weight operand is built)
    against an explicit flip + convolution; with a LeakyReLU' mask and accumulate."""
    ops = OPS.CudaOps(precision='fp32')
    w = rnd(5, 5, 4, 4, seed=1)
    dy, aux, acc0 = rnd(N, T, F, 4, seed=2), rnd(N, T, F, 4, seed=3), rnd(N, T, F, 4, seed=4)
    wt = torch.zeros(5, 5, 4, 4, dtype=torch.float64)
    REF.conv2d_flip(w, wt)
    for flags, kw in ((0, {}), (0, dict(aux=aux)), (OPS.EPI_ACCUM, {})):
        want = acc0.clone()
        REF.conv2d(dy, wt, want, flags=flags, **kw)
        got = dev(acc0)
        scratch = torch.zeros(5, 5, 4, 4, device='cuda')
        n0 = ops.launches
        ops.conv2d_dgrad(dev(dy), dev(w), got, scratch, flags=flags, **dict((k, dev(v)) for k, v in kw.items()))
        assert ops.launches == n0 + 1          # no flip launch
        close(got, want, 2e-5, ('conv2d_dgrad', flags, list(kw)))


def test_lse_column_window_and_add3(cuda_ops):
    """pctts_lse_loss_grad_cols over two column windows = pctts_lse_loss_grad over all columns (loss sums add up, gradients are the
    same elements); pctts_add3."""
    M, O = 777, 163
    pred, real, w = dev(rnd(M, O)), dev(rnd(M, O, seed=1)), dev(rnd(O, seed=2).abs())
    losses = torch.zeros(9, device='cuda')
    d_all, d_win = torch.zeros(M, O, device='cuda'), torch.full((M, O), 7.0, device='cuda')
    cuda_ops.lse_loss_grad(pred, real, w, d_all, 0.5, 0.25, losses, 0)
    cuda_ops.lse_loss_grad_cols(pred, real, w, d_win, 1, O - 1, 0.5, 0.25, losses, 1)
    cuda_ops.lse_loss_grad_cols(pred, real, w, d_win, 0, 1, 0.5, 0.25, losses, 2)
    torch.cuda.synchronize()
    assert torch.equal(d_all, d_win)
    assert abs(float(losses[0]) - float(losses[1] + losses[2])) < 1e-5 * abs(float(losses[0]))
    a, b, c = dev(rnd(1000, 16)), dev(rnd(1000, 16, seed=1)), dev(rnd(1000, 16, seed=2))
    out = torch.zeros(1000, 16, device='cuda')
    cuda_ops.add3(out, a, b, c)
    assert torch.equal(out, (a + b) + c)


def test_conv2d_strided_feature_views(cuda_ops):
    """First critic layer reads / last generator layer writes the spectral columns of a [B,T,O] tensor in place."""
    B, T, S, O = 2, 21, 17, 23
    feats = rnd(B, T, O)
    w = rnd(5, 5, 1, 4, seed=1)
    want = torch.zeros(B, T, S, 4, dtype=torch.float64)
    REF.conv2d(feats[:, :, 1:1 + S].unsqueeze(-1), w, want)
    got = torch.zeros(B, T, S, 4, device='cuda')
    cuda_ops.conv2d(dev(feats)[:, :, 1:1 + S].unsqueeze(-1), dev(w), got)
    close(got, want, 2e-5)
    w2 = rnd(5, 5, 4, 1, seed=2)
    outw = feats.clone()
    REF.conv2d(want, w2, outw[:, :, 1:1 + S].unsqueeze(-1), flags=OPS.EPI_ACCUM)
    outg = dev(feats)
    cuda_ops.conv2d(dev(want), dev(w2), outg[:, :, 1:1 + S].unsqueeze(-1), flags=OPS.EPI_ACCUM)
    close(outg, outw, 2e-5)


@pytest.mark.parametrize('rows,C', [(4000, 256), (1000 * 17, 4), (333, 2), (77, 300), (5, 8), (10 * 400 * 129, 4), (1031, 4),
                                    (9, 1028), (12000, 256)])
def test_batchnorm_forward_backward(cuda_ops, rows, C):
    x, dy = rnd(rows, C) * 2 + 0.3, rnd(rows, C, seed=1)
    gamma, beta = rnd(C, seed=2) + 1.5, rnd(C, seed=3) * 0.1
    mm, mv = rnd(C, seed=4), rnd(C, seed=5).abs() + 0.5
    ws = REF.bn_stats(x)
    wsd = cuda_ops.bn_stats(dev(x))
    close(wsd.sums, ws.sums, 1e-6, 'bn sums')
    mean, istd = torch.zeros(C, dtype=torch.float64), torch.zeros(C, dtype=torch.float64)
    mm_w, mv_w = mm.clone(), mv.clone()
    REF.bn_finalize(ws, rows, 1e-3, 0.99, mean, istd, mm_w, mv_w)
    meand, istdd, mmd, mvd = torch.zeros(C, device='cuda'), torch.zeros(C, device='cuda'), dev(mm), dev(mv)
    cuda_ops.bn_finalize(wsd, rows, 1e-3, 0.99, meand, istdd, mmd, mvd)
    for g_, w_ in ((meand, mean), (istdd, istd), (mmd, mm_w), (mvd, mv_w)):
        close(g_, w_, 1e-5, 'bn finalize')
    for use_var, second in ((0, istd), (1, mv)):
        y = torch.zeros(rows, C, dtype=torch.float64)
        REF.bn_apply(x, y, mean, second, gamma, beta, 1e-3, use_var, 1)
        yd = torch.zeros(rows, C, device='cuda')
        cuda_ops.bn_apply(dev(x), yd, dev(mean), dev(second), dev(gamma), dev(beta), 1e-3, use_var, 1)
        close(yd, y, 1e-5, 'bn apply')
    wsb = REF.bn_bwd_reduce(x, dy, mean, istd, gamma, beta, 1)
    dx, dg, db = torch.zeros(rows, C, dtype=torch.float64), torch.zeros(C, dtype=torch.float64), torch.zeros(C, dtype=torch.float64)
    REF.bn_bwd_apply(wsb, x, dy, dx, mean, istd, gamma, beta, rows, 1, dg, db, grad_scale=0.5)
    args = [dev(t) for t in (x, dy)]
    stat = [dev(t) for t in (mean, istd, gamma, beta)]
    wsbd = cuda_ops.bn_bwd_reduce(args[0], args[1], *stat, 1)
    close(wsbd.sums, wsb.sums, 2e-5, 'bn bwd sums')
    dxd, dgd, dbd = torch.zeros(rows, C, device='cuda'), torch.zeros(C, device='cuda'), torch.zeros(C, device='cuda')
    cuda_ops.bn_bwd_apply(wsbd, args[0], args[1], dxd, *stat, rows, 1, dgd, dbd, grad_scale=0.5)
    close(dxd, dx, 2e-5, 'bn dx')
    close(dgd, dg, 2e-5, 'bn dgamma')
    close(dbd, db, 2e-5, 'bn dbeta')
    cs = torch.zeros(C, dtype=torch.float64)
    REF.colsum(x, cs)
    csd = torch.zeros(C, device='cuda')
    cuda_ops.colsum(dev(x), csd)
    close(csd, cs, 1e-5, 'colsum')


# H in {32,64,128,256}: register-resident dataflow kernels (lstm_fast.cu); other H: generic shared-memory kernels (lstm.cu)
@pytest.mark.parametrize('B,T,H', [(10, 50, 256), (3, 17, 8), (2, 9, 2), (1, 1, 4), (5, 33, 24), (3, 400, 256), (1, 1, 256), (1, 2, 32),
                                   (5, 21, 32), (4, 30, 64), (7, 13, 128)])
def test_lstm_forward_and_bptt(cuda_ops, B, T, H):
    zx = [rnd(B, T, 4 * H, seed=d) for d in range(2)]
    U = [rnd(H, 4 * H, seed=2 + d) * (1.0 / np.sqrt(H)) for d in range(2)]
    hout = torch.zeros(B, T, 2 * H, dtype=torch.float64)
    gates = [torch.zeros(B, T, 4 * H, dtype=torch.float64) for _ in range(2)]
    cst = [torch.zeros(B, T, H, dtype=torch.float64) for _ in range(2)]
    REF.lstm_fwd(zx, U, hout, gates, cst)
    zxd, Ud = [dev(t) for t in zx], [dev(t) for t in U]
    houtd = torch.zeros(B, T, 2 * H, device='cuda')
    gatesd = [torch.zeros(B, T, 4 * H, device='cuda') for _ in range(2)]
    cstd = [torch.zeros(B, T, H, device='cuda') for _ in range(2)]
    cuda_ops.lstm_fwd(zxd, Ud, houtd, gatesd, cstd)
    close(houtd, hout, 1e-4, 'lstm h')
    for d in range(2):
        close(gatesd[d], gates[d], 1e-4, 'lstm gates')
        close(cstd[d], cst[d], 1e-4, 'lstm c')
    dh = rnd(B, T, 2 * H, seed=9)
    dz = [torch.zeros(B, T, 4 * H, dtype=torch.float64) for _ in range(2)]
    REF.lstm_bwd(dh, U, gates, cst, dz)
    dzd = [torch.zeros(B, T, 4 * H, device='cuda') for _ in range(2)]
    cuda_ops.lstm_bwd(dev(dh), Ud, [dev(t) for t in gates], [dev(t) for t in cst], dzd)
    for d in range(2):
        close(dzd[d], dz[d], 2e-4, 'lstm dz')
    # both kernel families agree (a negative cluster_size forces the generic kernels) and one direction alone works
    hg = torch.zeros(B, T, 2 * H, device='cuda')
    cuda_ops.lstm_fwd(zxd, Ud, hg, None, None, cluster_size=-8)
    close(hg, hout, 1e-4, 'lstm h generic')
    h1 = torch.zeros(B, T, H, device='cuda')
    cuda_ops.lstm_fwd(zxd[:1], Ud[:1], h1)
    close(h1, hout[:, :, :H], 1e-4, 'lstm single direction')


def test_gp_norms_sliced_demo_size(cuda_ops):
    """Per-sample squared gradient norm at the demo size (16 CTAs per sample, ticketed deterministic finish): against fp64,
    repeated launches (the ticket re-arms itself) bit-identical, ragged per-sample length."""
    for B, n in ((10, 400 * 129), (3, 4097)):
        g = rnd(B, n, seed=3) * 0.05
        want = (g * g).sum(1)
        gd, ssd = dev(g), torch.zeros(B, device='cuda')
        cuda_ops.gp_norms(gd, ssd)
        close(ssd, want, 1e-6, 'ss sliced')
        first = ssd.clone()
        for _ in range(3):
            ssd.zero_()
            cuda_ops.gp_norms(gd, ssd)
            assert torch.equal(ssd, first)


def test_wgan_gp_pieces(cuda_ops):
    B, T, O, S = 4, 31, 23, 17
    real, fake, alpha = rnd(B, T, O), rnd(B, T, O, seed=1), torch.rand(B, dtype=torch.float64)
    s0 = torch.zeros(3 * B, T, S, dtype=torch.float64)
    REF.critic_pack_inputs(real, fake, alpha, s0, S, 1)
    s0d = torch.zeros(3 * B, T, S, device='cuda')
    cuda_ops.critic_pack_inputs(dev(real), dev(fake), dev(alpha), s0d, S, 1)
    close(s0d, s0, 1e-6, 'pack')
    p = torch.zeros(30, device='cuda')
    cuda_ops.fill_segments(p, 10, -1.5, 15, 2.5, 5, 1.0)
    assert p.cpu().tolist() == [-1.5] * 10 + [2.5] * 15 + [1.0] * 5
    losses = torch.zeros(8, dtype=torch.float64)
    lossesd = torch.zeros(8, device='cuda')
    x = rnd(5000)
    REF.sum_scaled(x, 0.25, losses, 1)
    cuda_ops.sum_scaled(dev(x), 0.25, lossesd, 1)
    g = rnd(B, T * S, seed=3) * 0.05
    ss, r0 = torch.zeros(B, dtype=torch.float64), torch.zeros(B, T * S, dtype=torch.float64)
    REF.gp_norms(g, ss)
    REF.gp_scale(g, ss, r0, 10.0, 2 * B, losses, 2)
    ssd, r0d = torch.zeros(B, device='cuda'), torch.zeros(B, T * S, device='cuda')
    cuda_ops.gp_norms(dev(g), ssd)
    cuda_ops.gp_scale(dev(g), ssd, r0d, 10.0, 2 * B, lossesd, 2)
    close(ssd, ss, 1e-5, 'ss')
    close(r0d, r0, 1e-5, 'r0')
    pred, w = rnd(B, T, O, seed=4), torch.rand(O, dtype=torch.float64)
    dp = torch.zeros(B, T, O, dtype=torch.float64)
    REF.lse_loss_grad(pred, real, w, dp, 1.0 / (B * T * O), 0.5, losses, 3)
    dpd = torch.zeros(B, T, O, device='cuda')
    cuda_ops.lse_loss_grad(dev(pred), dev(real), dev(w), dpd, 1.0 / (B * T * O), 0.5, lossesd, 3)
    cuda_ops.lse_loss_grad(dev(pred), dev(real), dev(w), dpd, 1.0 / (B * T * O), 0.5, lossesd, 3)   # re-armed counter
    close(dpd, dp, 1e-5, 'lse grad')
    REF.loss_combine(losses, 4, 1, 1.0, 2, 10.0, 3, -2.0)
    cuda_ops.loss_combine(lossesd, 4, 1, 1.0, 2, 10.0, 3, -2.0)
    close(lossesd, losses, 1e-5, 'losses')


def test_adam_matches_keras_update(cuda_ops):
    n = 100003
    p, g, m, v = rnd(n), rnd(n, seed=1), rnd(n, seed=2) * 0.1, rnd(n, seed=3).abs() * 0.1
    step = torch.zeros(1, dtype=torch.int32)
    pd, gd, md, vd, stepd = dev(p), dev(g), dev(m), dev(v), torch.zeros(1, dtype=torch.int32, device='cuda')
    for _ in range(3):
        REF.adam(p, g, m, v, 1e-3, 0.5, 0.9, 1e-7, step)
        cuda_ops.adam(pd, gd, md, vd, 1e-3, 0.5, 0.9, 1e-7, stepd)
    assert int(stepd.item()) == 3
    close(pd, p, 1e-6, 'adam p')
    close(md, m, 1e-6, 'adam m')
    close(vd, v, 1e-6, 'adam v')


def test_small_movers_and_uniform(cuda_ops):
    a, b = rnd(1000), rnd(1000, seed=1)
    out = torch.zeros(1000, device='cuda')
    cuda_ops.add(out, dev(a), dev(b))
    close(out, a + b, 1e-6)
    big, y = rnd(50, 23), torch.rand(50, 23, dtype=torch.float64)
    dz = torch.zeros(50, 5, device='cuda')
    cuda_ops.sigmoid_bwd(dev(big)[:, 18:], dev(y)[:, 18:], dz)
    close(dz, big[:, 18:] * y[:, 18:] * (1 - y[:, 18:]), 1e-6)
    dst = dev(rnd(50, 7, seed=2))
    want = dst.cpu().double() + big[:, 3:10]
    cuda_ops.copy2d(dst, dev(big)[:, 3:10], accumulate=True)
    close(dst, want, 1e-6)
    src = rnd(3, 11, 5)
    pd = torch.full((3, 11 + 4 + 6, 5), 7.0, device='cuda')
    cuda_ops.pad_time(pd, dev(src), 4, 6)
    want = torch.zeros(3, 21, 5, dtype=torch.float64)
    want[:, 4:15] = src
    close(pd, want, 1e-7)
    u = torch.zeros(4096, device='cuda')
    counter = torch.zeros(1, dtype=torch.int64, device='cuda')
    cuda_ops.uniform(u, 123, counter)
    u1 = u.clone()
    cuda_ops.uniform(u, 123, counter)
    assert int(counter.item()) == 2
    assert float(u.min()) >= 0.0 and float(u.max()) < 1.0
    assert abs(float(u.mean()) - 0.5) < 0.03 and not torch.equal(u, u1)


def test_peer_allreduce_single_rank_protocol(cuda_ops):
    """csrc/peer.cu with world = 1 (the rank is its own peer): the publish / flag / wait / rank-ordered-sum protocol runs on one
    GPU -- values come back unchanged, the sequence number advances, both parity slots get used, no time-out is raised.  (The
    multi-rank exchange is checked against NCCL semantics by tools/dp_gpu_check.py on 2 GPUs.)"""
    import ctypes
    lib = cuda_ops.lib
    nmax = 64
    nbytes = lib.pctts_peer_allreduce_bytes(1, nmax)
    assert nbytes > 0 and lib.pctts_peer_allreduce_bytes(17, nmax) < 0
    buf = torch.zeros((nbytes + 7) // 8, dtype=torch.float64, device='cuda')
    bases = (ctypes.c_ulonglong * 1)(buf.data_ptr())
    err = torch.zeros(1, dtype=torch.int32, device='cuda')
    st = torch.cuda.current_stream().cuda_stream
    for k in range(5):
        x = torch.arange(37, dtype=torch.float64, device='cuda') * (k + 1) + 0.25
        want = x.clone()
        rc = lib.pctts_peer_allreduce_f64(x.data_ptr(), 37, ctypes.cast(bases, ctypes.c_void_p), 0, 1, nmax, err.data_ptr(), st)
        assert rc == 0
        torch.cuda.synchronize()
        assert torch.equal(x, want) and int(err) == 0
        assert int(buf.view(torch.int64)[0]) == k + 1          # sequence number
    assert lib.pctts_peer_allreduce_f64(x.data_ptr(), nmax + 1, ctypes.cast(bases, ctypes.c_void_p), 0, 1, nmax, err.data_ptr(), st) == -1
