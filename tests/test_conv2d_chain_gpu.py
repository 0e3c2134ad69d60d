"""
GPU parity of the multi-layer Conv2D chain kernel (csrc/conv2d_chain.cu, pctts_conv2d_chain) against a float64 restatement of
the same layers: the critic's BatchNorm-free stack (networks_critic.py:66-68) forward, its data-gradient chain (kernels flipped
on the fly, LeakyReLU' masks from the stored activations, the input gradient as a 4->1 last layer, optionally accumulated into a
strided view) and the in-place tangent sweep of the gradient penalty.  Shapes cover 1 / 2 / 4 / 8 / 16 CTAs per sample, frame
counts that do not divide by the cluster size, one frame, and frequency widths that are not multiples of 4.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

ALPHA = 0.3
EPI_LRELU, EPI_ACCUM = 2, 16


def conv_same(x, w):
    """x [N,T,F,Cin] float64, w [5,5,Cin,Cout] -> [N,T,F,Cout] ('same' cross-correlation, keras semantics)."""
    return F.conv2d(F.pad(x.permute(0, 3, 1, 2), (2, 2, 2, 2)), w.permute(3, 2, 0, 1)).permute(0, 2, 3, 1)


def lrelu(x):
    return torch.where(x > 0, x, ALPHA * x)


def dl(y):
    return torch.where(y > 0, torch.ones_like(y), torch.full_like(y, ALPHA))


def flip(w):
    """forward kernel [5,5,Cin,Cout] -> data-gradient kernel [5,5,Cout,Cin]"""
    return torch.flip(w, dims=(0, 1)).permute(0, 1, 3, 2).contiguous()


def rel(got, want):
    return float((got.double() - want).abs().max() / (want.abs().max() + 1e-30))


def make_weights(L, seed, scale=0.25):
    g = torch.Generator(device='cuda').manual_seed(seed)
    ws = [torch.randn(5, 5, 1 if i == 0 else 4, 4, device='cuda', generator=g) * (0.6 if i == 0 else scale) for i in range(L)]
    bs = [torch.randn(4, device='cuda', generator=g) * 0.1 for _ in range(L)]
    return ws, bs


SHAPES = [(3, 1, 129, 8), (2, 7, 129, 8), (2, 50, 129, 8), (2, 55, 129, 3), (3, 130, 129, 8), (1, 293, 129, 8), (3, 400, 129, 8),
          (1, 747, 129, 8), (2, 61, 33, 4), (2, 90, 65, 2), (2, 37, 30, 1),
          (1, 226, 227, 2), (1, 211, 243, 2), (1, 225, 227, 2)]   # wide maps: 16 CTAs, the last one owning a single frame / refused (skipped)


@pytest.mark.parametrize('N,T,Fq,L', SHAPES)
def test_forward_chain(cuda_ops, N, T, Fq, L):
    ops = cuda_ops
    if not ops.conv2d_chain_ok(T, Fq, 5, 5, 4):
        pytest.skip('shape not supported by the chain kernel')
    ws, bs = make_weights(L, 1)
    g = torch.Generator(device='cuda').manual_seed(2)
    x = torch.randn(N, T, Fq, 1, device='cuda', generator=g)
    ys = [torch.full((N, T, Fq, 4), float('nan'), device='cuda') for _ in range(L)]
    ops.conv2d_chain(x, [dict(w=ws[i], bias=bs[i], y=ys[i], flags=EPI_LRELU) for i in range(L)])
    torch.cuda.synchronize()
    cur = x.double()
    for i in range(L):
        cur = lrelu(conv_same(cur, ws[i].double()) + bs[i].double())
        assert torch.isfinite(ys[i]).all(), i
        assert rel(ys[i], cur) < 2e-5, (i, rel(ys[i], cur))
        cur = ys[i].double()          # layer by layer: the next layer's reference starts from what the kernel handed on


@pytest.mark.parametrize('N,T,Fq,L', [(2, 7, 129, 8), (3, 130, 129, 8), (2, 400, 129, 8), (2, 61, 33, 4), (1, 50, 129, 1)])
@pytest.mark.parametrize('accumulate', [False, True])
def test_data_gradient_chain_with_input_gradient(cuda_ops, N, T, Fq, L, accumulate):
    ops = cuda_ops
    if not ops.conv2d_chain_ok(T, Fq, 5, 5, 4):
        pytest.skip('shape not supported by the chain kernel')
    ws, _ = make_weights(L, 3)
    g = torch.Generator(device='cuda').manual_seed(4)
    acts = [torch.randn(N, T, Fq, 4, device='cuda', generator=g) for _ in range(L)]          # stored forward activations a_i
    da_top = torch.randn(N, T, Fq, 4, device='cuda', generator=g)
    das = [torch.full((N, T, Fq, 4), float('nan'), device='cuda') for _ in range(L - 1)] + [da_top]
    O = 7 + Fq
    dpred = torch.randn(N, T, O, device='cuda', generator=g)                                  # strided target: columns 3 .. 3+F
    keep = dpred.clone()
    gout = dpred[:, :, 3:3 + Fq].unsqueeze(-1)
    layers = [dict(w=ws[i], flip=True, aux=acts[i - 1], y=das[i - 1]) for i in range(L - 1, 0, -1)]
    layers.append(dict(w=ws[0], flip=True, y=gout, flags=EPI_ACCUM if accumulate else 0))
    ops.conv2d_chain(da_top, layers)
    torch.cuda.synchronize()
    cur = da_top.double()
    for i in range(L - 1, 0, -1):
        want = conv_same(cur, flip(ws[i]).double()) * dl(acts[i - 1].double())
        assert rel(das[i - 1], want) < 2e-5, (i, rel(das[i - 1], want))
        cur = das[i - 1].double()
    want = conv_same(cur, flip(ws[0]).double())[..., 0]
    if accumulate:
        want = want + keep[:, :, 3:3 + Fq].double()
    assert rel(dpred[:, :, 3:3 + Fq], want) < 2e-5
    assert torch.equal(dpred[:, :, :3], keep[:, :, :3]) and torch.equal(dpred[:, :, 3 + Fq:], keep[:, :, 3 + Fq:])


@pytest.mark.parametrize('N,T,Fq,L', [(2, 7, 129, 8), (3, 130, 129, 8), (2, 400, 129, 8), (2, 61, 33, 4)])
def test_in_place_tangent_chain_on_a_sample_slice(cuda_ops, N, T, Fq, L):
    """The sweep of the gradient penalty: t_i = conv(t_{i-1}, w_i) * LeakyReLU'(a_i) written over a_i, on the last third of a
    batched buffer (the interpolated applications) -- the other samples must stay untouched."""
    ops = cuda_ops
    if not ops.conv2d_chain_ok(T, Fq, 5, 5, 4):
        pytest.skip('shape not supported by the chain kernel')
    ws, _ = make_weights(L, 5)
    g = torch.Generator(device='cuda').manual_seed(6)
    full = [torch.randn(3 * N, T, Fq, 4, device='cuda', generator=g) for _ in range(L)]
    keep = [a.clone() for a in full]
    s0 = torch.randn(3 * N, T, Fq, device='cuda', generator=g)
    rows = slice(2 * N, 3 * N)
    ops.conv2d_chain(s0[rows].unsqueeze(-1), [dict(w=ws[i], aux=full[i][rows], y=full[i][rows]) for i in range(L)])
    torch.cuda.synchronize()
    cur = s0[rows].unsqueeze(-1).double()
    for i in range(L):
        want = conv_same(cur, ws[i].double()) * dl(keep[i][rows].double())
        assert rel(full[i][rows], want) < 2e-5, (i, rel(full[i][rows], want))
        assert torch.equal(full[i][:2 * N], keep[i][:2 * N])
        cur = full[i][rows].double()


def test_chain_is_deterministic_and_matches_the_per_layer_kernels(cuda_ops):
    """Run to run bit-identical; the 4->4 layers agree with the single-layer tensor-core kernel to fp32 round-off (the same
    MMA sequence per accumulator row), the whole stack to 1e-5."""
    ops = cuda_ops
    N, T, Fq, L = 3, 400, 129, 8
    if not ops.conv2d_chain_ok(T, Fq, 5, 5, 4):
        pytest.skip('shape not supported by the chain kernel')
    ws, bs = make_weights(L, 7)
    x = torch.randn(N, T, Fq, 1, device='cuda')
    ya = [torch.empty(N, T, Fq, 4, device='cuda') for _ in range(L)]
    yb = [torch.empty(N, T, Fq, 4, device='cuda') for _ in range(L)]
    for ys in (ya, yb):
        ops.conv2d_chain(x, [dict(w=ws[i], bias=bs[i], y=ys[i], flags=EPI_LRELU) for i in range(L)])
    torch.cuda.synchronize()
    for i in range(L):
        assert torch.equal(ya[i], yb[i]), i
    cur = x
    for i in range(L):
        y = torch.empty(N, T, Fq, 4, device='cuda')
        ops.conv2d(cur, ws[i], y, bias=bs[i], flags=EPI_LRELU)
        assert rel(ya[i], y.double()) < 1e-5, i
        cur = ya[i]
