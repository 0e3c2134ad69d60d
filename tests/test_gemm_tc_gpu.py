"""
GPU parity tests of the tcgen05 / TMEM / TMA GEMM family (percivaltts_b200/csrc/gemm_tc.cu) through the C ABI:
all four transpose combinations (K-major and MN-major UMMA descriptors), ragged sizes (TMA zero fill), every fused
epilogue, split-K, and the virtual-im2col Conv1D forward / weight gradient, against float64 torch on the CPU.

Tolerances, relative to the output scale: 'fp32' (three bf16 planes, 6 products per k-slice; the default) 3e-6 -- fp32
round-off level; 'fp32x3' (two planes, 3 products) 1e-4; 'bf16' (single product) 2e-2 (stated separately, BASELINE config #3).
"""
import pytest
import torch

from cpu_ops_double import CpuOpsDouble
from percivaltts_b200 import ops as OPS

pytestmark = pytest.mark.gpu

REF = CpuOpsDouble(torch.float64)
TOL6, TOL3, TOL1 = 3e-6, 1e-4, 2e-2


def rnd(*shape, seed=0):
    g = torch.Generator().manual_seed(seed + 31 * len(shape) + sum(shape))
    return torch.randn(*shape, generator=g, dtype=torch.float64)


def dev(t):
    return t.to(torch.float32).cuda()


def err(got, want):
    got = got.detach().cpu().double()
    return float((got - want).abs().max()) / (float(want.abs().max()) + 1e-30)


@pytest.fixture(scope='module')
def ops_bf16():
    return OPS.CudaOps(precision='bf16')


TC_CASES = [
    # M, N, K, transA, transB
    (128, 128, 64, False, True),      # one tile, one k-block, both operands K-major
    (128, 128, 64, False, False),     # B MN-major
    (128, 128, 64, True, True),       # A MN-major
    (128, 128, 64, True, False),      # both MN-major
    (300, 70, 45, False, False), (129, 64, 16, False, True), (64, 33, 257, True, False), (100, 24, 90, True, True),
    (1000, 256, 300, False, False), (4000, 256, 256, False, False), (4000, 256, 256, False, True),
    (515, 129, 256, False, False), (777, 516, 256, False, True),
]


@pytest.mark.parametrize('M,N,K,ta,tb', TC_CASES)
def test_gemm_tc_plain(cuda_ops, M, N, K, ta, tb):
    assert cuda_ops.nterms == 6
    A = rnd(K, M) if ta else rnd(M, K)
    B = rnd(N, K, seed=1) if tb else rnd(K, N, seed=1)
    want = torch.zeros(M, N, dtype=torch.float64)
    REF.gemm(A, B, want, transA=ta, transB=tb)
    got = torch.full((M, N), float('nan'), device='cuda')
    cuda_ops.gemm(dev(A), dev(B), got, transA=ta, transB=tb)
    assert err(got, want) < TOL6, (M, N, K, ta, tb, err(got, want))


@pytest.mark.parametrize('M,N,K,ta,tb', [(300, 70, 130, False, False), (1000, 256, 300, False, True)])
def test_gemm_tc_epilogues(cuda_ops, M, N, K, ta, tb):
    A = rnd(K, M) if ta else rnd(M, K)
    B = rnd(N, K, seed=1) if tb else rnd(K, N, seed=1)
    bias, aux, addmod = rnd(N, seed=2), rnd(M, N, seed=3), rnd(max(M // 3, 1), N, seed=4)
    for flags, kw in ((OPS.EPI_LRELU, dict(bias=bias)), (OPS.EPI_SIGMOID, dict(bias=bias)), (0, dict(aux=aux)),
                      (OPS.EPI_LRELU, dict(addmod=addmod)), (OPS.EPI_ACCUM, dict(bias=bias))):
        C0 = rnd(M, N, seed=5)
        want = C0.clone()
        REF.gemm(A, B, want, transA=ta, transB=tb, alpha=0.7, flags=flags, **kw)
        got = dev(C0)
        cuda_ops.gemm(dev(A), dev(B), got, transA=ta, transB=tb, alpha=0.7, flags=flags,
                      **dict((k, dev(v)) for k, v in kw.items()))
        assert err(got, want) < TOL6, (flags, list(kw), err(got, want))
    # aux aliasing C (tangent sweep) and a strided C view
    C0 = rnd(M, N + 5, seed=6)
    want = C0.clone()
    REF.gemm(A, B, want[:, 2:2 + N], transA=ta, transB=tb, aux=want[:, 2:2 + N].clone())
    got = dev(C0)
    cuda_ops.gemm(dev(A), dev(B), got[:, 2:2 + N], transA=ta, transB=tb, aux=got[:, 2:2 + N])
    assert err(got, want) < TOL6


def test_gemm_tc_splitk_weight_gradients(cuda_ops):
    """dW = X^T dZ with a long reduction and few output tiles: split-K partials + deterministic reduce, with accumulate."""
    for M, N, K in ((256, 256, 4000), (516, 256, 12000), (256, 1024, 4000), (256, 33, 4000)):
        X, dZ = rnd(K, M), rnd(K, N, seed=1)
        assert cuda_ops.lib.pctts_gemm_tc_splits(M, N, K) > 1
        C0 = rnd(M, N, seed=2)
        for flags in (0, OPS.EPI_ACCUM):
            want = C0.clone()
            REF.gemm(X, dZ, want, transA=True, flags=flags)
            got = dev(C0)
            cuda_ops.gemm(dev(X), dev(dZ), got, transA=True, flags=flags)
            assert err(got, want) < TOL6, (M, N, K, flags, err(got, want))
        again = dev(C0)
        cuda_ops.gemm(dev(X), dev(dZ), again, transA=True, flags=OPS.EPI_ACCUM)
        assert torch.equal(again, got), 'split-K reduction must be deterministic'


def test_gemm_tc_tail_split_of_the_last_wave(cuda_ops):
    """More output tiles than SMs (critic post FC on the batched [real | fake | interp] rows: 188 tiles): the m-tiles of the
    partly filled last wave are cut along K and summed by the deterministic reduce, which then also carries the epilogue."""
    for M, N, K in ((12000, 256, 772), (11901, 250, 800)):
        assert cuda_ops.lib.pctts_gemm_tc_tail_workspace(M, N, K) > 0
        A, B = rnd(M, K), rnd(K, N, seed=1)
        bias, aux = rnd(N, seed=2), rnd(M, N, seed=3)
        for flags, kw in ((0, {}), (OPS.EPI_LRELU, dict(bias=bias)), (0, dict(aux=aux)), (OPS.EPI_ACCUM, dict(bias=bias))):
            C0 = rnd(M, N, seed=5)
            want = C0.clone()
            REF.gemm(A, B, want, alpha=0.7, flags=flags, **kw)
            got = dev(C0)
            cuda_ops.gemm(dev(A), dev(B), got, alpha=0.7, flags=flags, **dict((k, dev(v)) for k, v in kw.items()))
            assert err(got, want) < TOL6, (M, N, K, flags, list(kw), err(got, want))
            again = dev(C0)
            cuda_ops.gemm(dev(A), dev(B), again, alpha=0.7, flags=flags, **dict((k, dev(v)) for k, v in kw.items()))
            assert torch.equal(again, got)
    assert cuda_ops.lib.pctts_conv1d_tc_wgrad_workspace(10, 400, 21, 425, 256) > 0        # 168 tiles of 70 k-blocks
    assert cuda_ops.lib.pctts_gemm_tc_tail_workspace(4000, 256, 256) == 0                 # 64 tiles: nothing to cut


def test_gemm_tc_sliced_3d_operands(cuda_ops):
    """LSTM recurrent-kernel gradient pattern: sliced [B, T-1, H] views of [B, T, 2H] / [B, T, 4H] buffers."""
    Bn, T, H = 3, 50, 64
    h, dzl = rnd(Bn, T, 2 * H, seed=3), rnd(Bn, T, 4 * H, seed=4)
    want = torch.zeros(H, 4 * H, dtype=torch.float64)
    REF.gemm(h[:, 0:T - 1, H:], dzl[:, 1:T], want, transA=True)
    got = torch.zeros(H, 4 * H, device='cuda')
    cuda_ops.gemm(dev(h)[:, 0:T - 1, H:], dev(dzl)[:, 1:T], got, transA=True)
    assert err(got, want) < TOL6


@pytest.mark.parametrize('Bn,T,C,W,H', [(3, 29, 11, 5, 16), (2, 200, 70, 3, 64), (10, 400, 425, 21, 256), (2, 130, 425, 21, 256)])
def test_conv1d_tc_forward_and_wgrad(cuda_ops, Bn, T, C, W, H):
    left = (W - 1) // 2
    xp = torch.zeros(Bn, T + W - 1, C, dtype=torch.float64)
    xp[:, left:left + T] = rnd(Bn, T, C)
    k = rnd(W, C, H, seed=1) / (W * C) ** 0.5
    bias = rnd(H, seed=2)
    want = torch.zeros(Bn * T, H, dtype=torch.float64)
    REF.conv1d(xp, k, want, T, bias=bias, flags=OPS.EPI_LRELU)
    got = torch.full((Bn * T, H), float('nan'), device='cuda')
    xpd = dev(xp)
    cuda_ops.conv1d(xpd, dev(k), got, T, bias=dev(bias), flags=OPS.EPI_LRELU)
    assert err(got, want) < TOL6, ('fwd', err(got, want))
    dz = rnd(Bn * T, H, seed=3)
    wantw = torch.zeros(W, C, H, dtype=torch.float64)
    REF.conv1d_wgrad(xp, dz, wantw, T)
    gotw = torch.full((W, C, H), float('nan'), device='cuda')
    cuda_ops.conv1d_wgrad(xpd, dev(dz), gotw, T)
    assert err(gotw, wantw) < TOL6, ('wgrad', err(gotw, wantw))


def test_gemm_tc_fp32x3_mode():
    ops3 = OPS.CudaOps(precision='fp32x3')
    M, N, K = 1000, 256, 300
    A, B = rnd(M, K), rnd(K, N, seed=1)
    want = torch.zeros(M, N, dtype=torch.float64)
    REF.gemm(A, B, want)
    got = torch.zeros(M, N, device='cuda')
    ops3.gemm(dev(A), dev(B), got)
    e = err(got, want)
    assert 3e-7 < e < TOL3, e


def test_gemm_tc_bf16_mode(ops_bf16):
    M, N, K = 1000, 256, 300
    A, B = rnd(M, K), rnd(K, N, seed=1)
    want = torch.zeros(M, N, dtype=torch.float64)
    REF.gemm(A, B, want)
    got = torch.zeros(M, N, device='cuda')
    ops_bf16.gemm(dev(A), dev(B), got)
    e = err(got, want)
    assert 1e-5 < e < TOL1, e          # really the single-product path, and within the stated bf16 tolerance


def test_shadow_freshness_inside_step_scope(cuda_ops):
    """Inside a step scope a shadow is reused until one of the library's kernels overwrites its source."""
    M, N, K = 256, 64, 128
    A, B = dev(rnd(M, K)), dev(rnd(K, N, seed=1))
    C = torch.zeros(M, N, device='cuda')
    with cuda_ops.step_scope():
        n0 = cuda_ops.launches
        cuda_ops.gemm(A, B, C)
        first = cuda_ops.launches - n0
        n0 = cuda_ops.launches
        cuda_ops.gemm(A, B, C)
        assert cuda_ops.launches - n0 == first - 2          # both operands still fresh: no re-split
        cuda_ops.fill_segments(A.view(-1), A.numel(), 1.0)     # a writer invalidates the shadow of A
        n0 = cuda_ops.launches
        cuda_ops.gemm(A, B, C)
        assert cuda_ops.launches - n0 == first - 1
        want = torch.ones(M, K, dtype=torch.float64) @ B.cpu().double()
        assert err(C, want) < TOL6


def test_parameter_ranges_die_with_their_buffer():
    """A registered parameter buffer that has been freed must not make later tensors at the same address 'persistent'
    (their bf16 operand shadows would never be refreshed): regression test for a stale-range aliasing bug."""
    from percivaltts_b200.ops import CudaOps
    ops = CudaOps()
    flat = ops.zeros(64 * 64)
    ops.register_persistent(flat)
    ptr = flat.data_ptr()
    del flat
    a = ops.empty(64, 64)
    if a.data_ptr() != ptr:
        pytest.skip('allocator did not hand the freed block back')
    b = torch.randn(64, 64, device='cuda')
    c = ops.empty(64, 64)
    for _ in range(2):
        a.normal_()
        ops.gemm(a, b, c)
        want = a.double() @ b.double()
        assert float((c.double() - want).abs().max()) < 1e-4 * float(want.abs().max())
