"""
GPU parity of the tcgen05 implicit-GEMM Conv2D (csrc/conv2d_tc.cu, pctts_conv2d_tc_fwd) -- reference ops
networktts.py:122-126, modeltts_common.py:98, networks_critic.py:66-68 -- against the float64 CPU definition
(tests/cpu_ops_double.py) and against the exact-fp32 CUDA-core kernel, per precision mode.
Tolerances: fp32 (3 bf16 planes, products of order <= 2) 2e-6; fp32x3 (2 planes) 1e-4; bf16 (1 plane) 2e-2 of max|y|.
"""
import pytest
import torch

from cpu_ops_double import CpuOpsDouble
from percivaltts_b200 import ops as OPS

pytestmark = pytest.mark.gpu
REF = CpuOpsDouble(torch.float64)
TOL = {'fp32': 2e-6, 'fp32x3': 1e-4, 'bf16': 2e-2}


def rnd(*shape, seed=0):
    g = torch.Generator().manual_seed(seed + 31 * len(shape) + sum(shape))
    return torch.randn(*shape, generator=g, dtype=torch.float64)


def dev(t):
    return t.to(torch.float32).cuda()


def relerr(got, want):
    return float((got.detach().cpu().double() - want).abs().max()) / (float(want.abs().max()) + 1e-30)


@pytest.fixture(scope='module', params=['fp32', 'fp32x3', 'bf16'])
def tc_ops(request):
    if not torch.cuda.is_available():
        pytest.skip('no CUDA device')
    o = OPS.CudaOps(precision=request.param)
    assert o.conv2d_tc
    return o


# F covers: the demo width 129 (one valid pixel in the last group), multiples of 4 +- 1, tiny maps, T not a multiple of the tile
@pytest.mark.parametrize('N,T,F', [(2, 37, 129), (1, 5, 17), (3, 16, 65), (1, 1, 1), (2, 9, 2), (1, 23, 130), (5, 61, 33), (1, 400, 129)])
def test_conv2d_tc_all_epilogues(tc_ops, N, T, F):
    assert tc_ops.lib.pctts_conv2d_tc_supported(F, 4, 4, 5, 5, 2, 2) == 1
    tol = TOL[tc_ops.precision]
    x, w, b = rnd(N, T, F, 4), rnd(5, 5, 4, 4, seed=1) * 0.2, rnd(4, seed=2)
    aux = rnd(N, T, F, 4, seed=3)
    n0 = tc_ops.lib.pctts_launch_count()
    for flags, kw in ((0, {}), (OPS.EPI_LRELU, dict(bias=b)), (0, dict(aux=aux)), (OPS.EPI_ACCUM, {}), (OPS.EPI_ACCUM, dict(bias=b, aux=aux))):
        y0 = rnd(N, T, F, 4, seed=4)
        want = y0.clone()
        REF.conv2d(x, w, want, flags=flags, **kw)
        got = dev(y0)
        tc_ops.conv2d(dev(x), dev(w), got, flags=flags, **dict((k, dev(v)) for k, v in kw.items()))
        assert relerr(got, want) < tol, (flags, list(kw), relerr(got, want))
    # in-place tangent form: aux aliases y (Critic.tangent_forward)
    want = aux.clone()
    REF.conv2d(x, w, want, aux=want)
    got = dev(aux)
    tc_ops.conv2d(dev(x), dev(w), got, aux=got)
    assert relerr(got, want) < tol
    assert tc_ops.lib.pctts_launch_count() - n0 == 6


def test_conv2d_tc_matches_cuda_core_kernel_at_demo_size(tc_ops, cuda_ops_simt):
    """3B = 30 samples x 400 frames x 129 bins (the batched critic map): tensor-core path vs the exact fp32 kernel,
    data gradient form (flipped kernel + LeakyReLU' mask) and the sample-strided view of the last third of the batch."""
    tol = TOL[tc_ops.precision]
    x, w = dev(rnd(30, 400, 129, 4)), dev(rnd(5, 5, 4, 4, seed=1) * 0.2)
    aux = dev(rnd(30, 400, 129, 4, seed=3))
    wt = torch.empty(5, 5, 4, 4, device='cuda')
    tc_ops.conv2d_flip(w, wt)
    want, got = torch.empty_like(x), torch.empty_like(x)
    cuda_ops_simt.conv2d(x, wt, want, aux=aux)
    tc_ops.conv2d(x, wt, got, aux=aux)
    assert relerr(got, want.cpu().double()) < tol
    rows = slice(20, 30)
    want2, got2 = torch.zeros(10, 400, 129, 4, device='cuda'), torch.zeros(10, 400, 129, 4, device='cuda')
    cuda_ops_simt.conv2d(x[rows], w, want2)
    tc_ops.conv2d(x[rows], w, got2)
    assert relerr(got2, want2.cpu().double()) < tol
    # linearity in the input (size-independent property): conv(a x1 + x2) = a conv(x1) + conv(x2)
    x2 = dev(rnd(30, 400, 129, 4, seed=9))
    y1, y2, y12 = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
    tc_ops.conv2d(x, w, y1)
    tc_ops.conv2d(x2, w, y2)
    tc_ops.conv2d(0.5 * x + x2, w, y12)
    assert relerr(y12, (0.5 * y1 + y2).cpu().double()) < max(tol, 1e-5) * 4


def test_conv2d_tc_unsupported_shapes_take_the_cuda_core_kernel(tc_ops):
    for args in ((129, 1, 4, 5, 5, 2, 2), (129, 4, 1, 5, 5, 2, 2), (129, 4, 4, 3, 3, 1, 1), (129, 4, 4, 5, 5, 1, 2)):
        assert tc_ops.lib.pctts_conv2d_tc_supported(*args) == 0
    x, w = rnd(2, 11, 17, 1), rnd(5, 5, 1, 4, seed=1)
    want = torch.zeros(2, 11, 17, 4, dtype=torch.float64)
    REF.conv2d(x, w, want)
    got = torch.zeros(2, 11, 17, 4, device='cuda')
    tc_ops.conv2d(dev(x), dev(w), got)
    assert relerr(got, want) < 2e-5


@pytest.mark.parametrize('N,T,F', [(10, 400, 129), (3, 250, 65), (2, 40, 33)])   # the last is too small: falls back to bn_stats
def test_fused_batchnorm_statistics(cuda_ops, N, T, F):
    """BatchNorm sums [sum y, sum y^2] from the Conv2D epilogue (pctts_conv2d_tc_fwd_stats) equal the sums of a separate pass
    over the stored output (pctts_bn_stats) to fp64 round-off, and the stored output is unchanged by the fusion."""
    x, w = dev(rnd(N, T, F, 4)), dev(rnd(5, 5, 4, 4, seed=1) * 0.2)
    y0, y1 = torch.empty(N, T, F, 4, device='cuda'), torch.empty(N, T, F, 4, device='cuda')
    cuda_ops.conv2d(x, w, y0)
    want = cuda_ops.bn_stats(y0.view(-1, 4)).sums.clone()
    ws = cuda_ops.conv2d(x, w, y1, stats=True)
    assert torch.equal(y0, y1)
    got = ws.sums
    assert float(((got - want).abs() / want.abs().clamp_min(1e-30)).max()) < 1e-12
    yd = y0.double().view(-1, 4)
    ref = torch.stack([yd.sum(0), (yd * yd).sum(0)])
    assert float(((got - ref).abs() / ref.abs().clamp_min(1e-30)).max()) < 1e-10


@pytest.mark.parametrize('N,T,F', [(10, 400, 129), (3, 250, 65), (2, 40, 33)])
def test_single_launch_second_stage_and_finalize(cuda_ops, N, T, F):
    """Single replica: statistics second stage + finalize in one launch (pctts_bn_finalize_partials) from the partials of the
    Conv2D epilogue or of pctts_bn_stats_partial -- mean / invstd / moving averages / sums bit-identical to the two-launch path."""
    x, w = dev(rnd(N, T, F, 4)), dev(rnd(5, 5, 4, 4, seed=1) * 0.2)
    y = torch.empty(N, T, F, 4, device='cuda')
    rows = N * T * F

    def finalize(ws):
        mean, invstd = torch.empty(4, device='cuda'), torch.empty(4, device='cuda')
        mm, mv = torch.full((4,), 0.25, device='cuda'), torch.full((4,), 1.5, device='cuda')
        cuda_ops.bn_finalize(ws, rows, 1e-3, 0.99, mean, invstd, mm, mv)
        return mean, invstd, mm, mv, ws.sums.clone()

    ref = finalize(cuda_ops.conv2d(x, w, y, stats=True))
    ws = cuda_ops.conv2d(x, w, y, stats=True, stats_final=False)
    assert ws.pending > 0
    got = finalize(ws)
    assert ws.pending == 0 and all(torch.equal(a, b) for a, b in zip(ref, got))
    ref = finalize(cuda_ops.bn_stats(y.view(-1, 4)))
    got = finalize(cuda_ops.bn_stats(y.view(-1, 4), final=False))
    assert all(torch.equal(a, b) for a, b in zip(ref, got))
    z = dev(rnd(4000, 256, seed=7))                    # a dense layer's pre-activations
    refz = cuda_ops.bn_stats(z).sums.clone()
    wz = cuda_ops.bn_stats(z, final=False)
    mean, invstd = torch.empty(256, device='cuda'), torch.empty(256, device='cuda')
    cuda_ops.bn_finalize(wz, 4000, 1e-3, 0.99, mean, invstd)
    assert torch.equal(wz.sums, refz)
    zd = z.double()
    assert float((mean.double() - zd.mean(0)).abs().max()) < 1e-6
    assert float((invstd.double() - 1.0 / torch.sqrt(zd.var(0, unbiased=False) + 1e-3)).abs().max()) < 1e-5


@pytest.mark.parametrize('N,T,F', [(10, 400, 129), (2, 37, 30)])
def test_fused_input_batchnorm(cuda_ops, N, T, F):
    """LeakyReLU(BN(x)) formed on load by the convolution (in_bn) equals the convolution of the materialised activation map
    bit for bit (same arithmetic, same order), zero padding included."""
    x, w = dev(rnd(N, T, F, 4)), dev(rnd(5, 5, 4, 4, seed=1) * 0.2)
    mean, invstd = dev(rnd(4, seed=2) * 0.3), dev(rnd(4, seed=3).abs() + 0.5)
    gamma, beta = dev(rnd(4, seed=4)), dev(rnd(4, seed=5) + 0.7)      # beta != 0: a transformed zero pad would show
    act = torch.empty_like(x)
    cuda_ops.bn_apply(x.view(-1, 4), act.view(-1, 4), mean, invstd, gamma, beta, 0.0, 0, 1)
    y0, y1 = torch.empty_like(x), torch.empty_like(x)
    cuda_ops.conv2d(act, w, y0)
    scratch = torch.full_like(x, float('nan'))
    ws = cuda_ops.conv2d(x, w, y1, stats=True, in_bn=(mean, invstd, gamma, beta, scratch))
    assert torch.equal(y0, y1)
    assert bool(torch.isnan(scratch).all())          # nothing was materialised
    want = cuda_ops.bn_stats(y0.view(-1, 4)).sums.clone()
    assert float(((ws.sums - want).abs() / want.abs().clamp_min(1e-30)).max()) < 1e-12
