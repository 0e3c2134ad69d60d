"""
Guard-band tests of the round-2 kernels: compute-sanitizer is closed on this GPU pool, so every output (and scratch) buffer of a
call is carved out of a larger allocation filled with a sentinel, and the bands on either side must still hold the sentinel
afterwards -- an out-of-bounds WRITE of the chain kernel, the one-launch stack weight gradient (producer warp + bulk copies), the
plane-writing BatchNorm kernels, the multi-matrix split or the fused Adam would show up here.  (Out-of-bounds reads show up in the
parity tests, whose inputs sit between NaN-filled neighbours here as well.)  Shapes are ragged on purpose.
"""
import ctypes

import pytest
import torch

from percivaltts_b200 import ops as OPS

pytestmark = pytest.mark.gpu
SENT = -12345.678


class Arena(object):
    """One big sentinel-filled allocation; carve() hands out views separated by guard bands."""

    def __init__(self, nfloats, band=4096):
        self.buf = torch.full((nfloats,), SENT, device='cuda')
        self.pos, self.band, self.used = band, band, []

    def carve(self, *shape, fill=None):
        n = 1
        for s in shape:
            n *= s
        start = (self.pos + 3) // 4 * 4                      # 16-byte aligned views
        v = self.buf[start:start + n].view(*shape)
        self.used.append((start, start + n))
        self.pos = start + n + self.band
        assert self.pos <= self.buf.numel(), 'arena too small'
        if fill is not None:
            v.copy_(fill)
        return v

    def check(self, what):
        torch.cuda.synchronize()
        mask = torch.ones(self.buf.numel(), dtype=torch.bool, device='cuda')
        for a, b in self.used:
            mask[a:b] = False
        bad = (self.buf[mask] != SENT).nonzero()
        assert bad.numel() == 0, (what, 'guard band overwritten', int(bad.numel()))


@pytest.mark.parametrize('N,T,F,L', [(2, 55, 129, 8), (1, 7, 33, 3), (3, 130, 65, 5), (2, 61, 30, 2)])
def test_chain_and_stack_weight_gradient_stay_inside_their_buffers(N, T, F, L):
    ops = OPS.CudaOps(precision='fp32')
    if not ops.conv2d_chain_ok(T, F, 5, 5, 4):
        pytest.skip('shape not supported by the chain kernel')
    g = torch.Generator(device='cuda').manual_seed(1)
    rnd = lambda *s: torch.randn(*s, device='cuda', generator=g)
    ar = Arena(2 * (L + 3) * N * T * F * 4 + 200000)
    x = ar.carve(N, T, F, 1, fill=rnd(N, T, F, 1))
    ys = [ar.carve(N, T, F, 4) for _ in range(L)]
    ws = [rnd(5, 5, 1 if i == 0 else 4, 4) * 0.25 for i in range(L)]
    bs = [rnd(4) * 0.1 for _ in range(L)]
    ops.conv2d_chain(x, [dict(w=ws[i], bias=bs[i], y=ys[i], flags=OPS.EPI_LRELU) for i in range(L)])
    ar.check('chain forward')
    assert all(bool(torch.isfinite(y).all()) for y in ys)
    # data-gradient chain with masks, input gradient into a strided feature view
    das = [ar.carve(N, T, F, 4) for _ in range(L)]
    das[L - 1].copy_(rnd(N, T, F, 4))
    feats = ar.carve(N, T, F + 3)
    feats.zero_()
    layers = [dict(w=ws[i], flip=True, aux=ys[i - 1], y=das[i - 1]) for i in range(L - 1, 0, -1)]
    layers.append(dict(w=ws[0], flip=True, y=feats[:, :, 1:1 + F].unsqueeze(-1)))
    ops.conv2d_chain(das[L - 1], layers)
    ar.check('chain data gradient')
    assert float(feats[:, :, 0].abs().max()) == 0.0 and float(feats[:, :, 1 + F:].abs().max()) == 0.0
    # the stack's weight + bias gradients in one launch (workspace inside the arena too)
    dws = [ar.carve(5, 5, 1 if i == 0 else 4, 4) for i in range(L)]
    dbs = [ar.carve(4) for _ in range(L)]
    ws_multi = ar.carve(int(ops.lib.pctts_conv2d_wgrad_multi_workspace(L)))
    arr = (OPS.WgradLayerStruct * L)()
    for i in range(L):
        xi = x if i == 0 else ys[i - 1]
        arr[i] = OPS.WgradLayerStruct(xi.data_ptr(), das[i].data_ptr(), dws[i].data_ptr(), dbs[i].data_ptr(), xi.stride(0), xi.stride(1),
                                      das[i].stride(0), das[i].stride(1), int(xi.shape[3]))
    OPS._lib.check(ops.lib.pctts_conv2d_wgrad_multi(ctypes.cast(arr, ctypes.c_void_p), L, max(N - 1, 1), ws_multi.data_ptr(), N, T, F, 1.0, 0,
                                                    ops._stream()), 'wgrad multi')
    ar.check('stack weight gradient')
    assert all(bool(torch.isfinite(d).all()) for d in dws + dbs)


@pytest.mark.parametrize('rows,C', [(517, 64), (33, 8), (4001, 256)])
def test_plane_writing_batchnorm_and_multi_split_stay_inside_their_buffers(rows, C):
    ops = OPS.CudaOps(precision='fp32')
    g = torch.Generator(device='cuda').manual_seed(2)
    rnd = lambda *s: torch.randn(*s, device='cuda', generator=g)
    ar = Arena(12 * rows * C + 400000)
    x, dy = rnd(rows, C), rnd(rows, C)
    gamma, beta, mean, istd = rnd(C) + 1.5, rnd(C) * 0.1, rnd(C) * 0.1, rnd(C).abs() + 0.5
    y, dx = ar.carve(rows, C), ar.carve(rows, C)
    planes = ar.carve(3 * rows * C // 2 + 8)                 # 3 bf16 planes of [rows, C] = 1.5 floats per element
    rc = ops.lib.pctts_bn_apply_planes(x.data_ptr(), y.data_ptr(), mean.data_ptr(), istd.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                       rows, C, 1e-3, 0, 1, planes.data_ptr(), C, rows * C, 3, ops._stream())
    assert rc == 0
    ar.check('bn_apply_planes')
    ws = ops.bn_bwd_reduce(x, dy, mean, istd, gamma, beta, 1)
    rc = ops.lib.pctts_bn_bwd_apply_planes(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), mean.data_ptr(), istd.data_ptr(), gamma.data_ptr(),
                                           beta.data_ptr(), rows, C, float(rows), 1, ws.ptr, planes.data_ptr(), C, rows * C, 3, ops._stream())
    assert rc == 0
    ar.check('bn_bwd_apply_planes')
    # multi-matrix split: three matrices of different shapes into plane buffers inside the arena
    mats = [rnd(rows, C), rnd(37, 24), rnd(5, 8)]
    dsts = [ar.carve(3 * m.shape[0] * ((m.shape[1] + 7) // 8 * 8) // 2 + 8) for m in mats]
    jobs = (OPS.SplitJobStruct * 3)(*[OPS.SplitJobStruct(m.data_ptr(), d.data_ptr(), m.stride(0), m.shape[0], (m.shape[1] + 7) // 8 * 8,
                                                         m.shape[0] * ((m.shape[1] + 7) // 8 * 8), m.shape[1]) for m, d in zip(mats, dsts)])
    OPS._lib.check(ops.lib.pctts_split_bf16_multi(ctypes.cast(jobs, ctypes.c_void_p), 3, 3, ops._stream()), 'split multi')
    ar.check('split_bf16_multi')


def test_fused_adam_single_rank_stays_inside_its_buffers():
    ops = OPS.CudaOps(precision='fp32')
    n = 100003
    npad = (n + 3) // 4 * 4
    ar = Arena(5 * npad + 100000)
    g = torch.Generator(device='cuda').manual_seed(3)
    grad, p, m, v = ar.carve(npad), ar.carve(n), ar.carve(n), ar.carve(n)
    grad.zero_()
    grad[:n].copy_(torch.randn(n, device='cuda', generator=g) * 0.1)
    p.copy_(torch.randn(n, device='cuda', generator=g))
    m.zero_()
    v.zero_()
    step = torch.zeros(1, dtype=torch.int32, device='cuda')
    ctl = torch.zeros(int(ops.lib.pctts_grad_sync_ctl_bytes()) // 8, dtype=torch.int64, device='cuda')
    err = torch.zeros(1, dtype=torch.int32, device='cuda')
    gb, cb = (ctypes.c_ulonglong * 1)(grad.data_ptr()), (ctypes.c_ulonglong * 1)(ctl.data_ptr())
    OPS._lib.check(ops.lib.pctts_grad_sync_adam(ctypes.cast(gb, ctypes.c_void_p), ctypes.cast(cb, ctypes.c_void_p), 0, 1, p.data_ptr(),
                                                m.data_ptr(), v.data_ptr(), n, npad, 1e-3, 0.5, 0.9, 1e-7, step.data_ptr(), None, 0,
                                                err.data_ptr(), ops._stream()), 'grad_sync_adam')
    ar.check('grad_sync_adam')
    assert int(step) == 1 and bool(torch.isfinite(p).all())
