"""
sync-BatchNorm reductions with the replica exchange inside the launch (csrc/peer.cu: bn_sync_kernel, pctts_bn_sync_finalize /
pctts_bn_sync_bwd): bit-identity with the separate launches (second stage, exchange, finalize / dgamma-dbeta) on one replica, and
the sliced publish / flag / wait / rank-ordered-sum protocol exercised by TWO ranks on ONE GPU (two symmetric buffers, two
streams), as the single-GPU stand-in for the 2..8-rank runs of tests/test_dp_gpu.py.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

EPS, MOM = 1e-3, 0.99


class _Local(object):
    world, rank, bn_handle = 1, 0, None


@pytest.mark.parametrize('rows,C', [(4000, 256), (516000, 4), (300, 5), (77, 513)])
def test_fused_second_stage_is_bit_identical_on_one_replica(cuda_ops, rows, C):
    o = cuda_ops
    g = torch.Generator(device='cuda').manual_seed(1)
    x = torch.randn(rows, C, device='cuda', generator=g) * 1.5 + 0.3
    dy = torch.randn(rows, C, device='cuda', generator=g)
    gam, bet = torch.rand(C, device='cuda', generator=g) + 0.5, torch.randn(C, device='cuda', generator=g) * 0.1
    mean_a, inv_a, mm_a, mv_a = (torch.zeros(C, device='cuda') for _ in range(4))
    mean_b, inv_b, mm_b, mv_b = (torch.zeros(C, device='cuda') for _ in range(4))
    mv_a.fill_(1.0), mv_b.fill_(1.0)
    ws = o.bn_stats(x, final=True)
    o.bn_finalize(ws, rows, EPS, MOM, mean_a, inv_a, mm_a, mv_a)
    sums_a = ws.sums.clone()
    ws = o.bn_stats(x, final=False)
    o.bn_sync_finalize(ws, rows, EPS, MOM, mean_b, inv_b, mm_b, mv_b, _Local())
    torch.cuda.synchronize()
    for a, b in ((mean_a, mean_b), (inv_a, inv_b), (mm_a, mm_b), (mv_a, mv_b), (sums_a, ws.sums)):
        assert torch.equal(a, b)
    # backward: sums + dgamma / dbeta
    dg_a, db_a, dx_a = torch.zeros(C, device='cuda'), torch.zeros(C, device='cuda'), torch.empty_like(x)
    dg_b, db_b, dx_b = torch.zeros(C, device='cuda'), torch.zeros(C, device='cuda'), torch.empty_like(x)
    ws = o.bn_bwd_reduce(x, dy, mean_a, inv_a, gam, bet, 1)
    o.bn_bwd_apply(ws, x, dy, dx_a, mean_a, inv_a, gam, bet, rows, 1, dg_a, db_a, grad_scale=0.5)
    ws = o.bn_bwd_reduce(x, dy, mean_a, inv_a, gam, bet, 1, final=False)
    o.bn_sync_bwd(ws, dg_b, db_b, 0.5, False, _Local())
    o.bn_bwd_apply(ws, x, dy, dx_b, mean_a, inv_a, gam, bet, rows, 1, None, None)
    torch.cuda.synchronize()
    assert torch.equal(dg_a, dg_b) and torch.equal(db_a, db_b) and torch.equal(dx_a, dx_b)


@pytest.mark.parametrize('C', [4, 256, 100])
def test_two_ranks_on_one_gpu_exchange_in_rank_order(cuda_ops, C):
    """Two 'ranks' = two symmetric buffers and two streams on this GPU; every call pairs a launch per rank.  The finalized
    statistics must be those of the rank-ordered sum of the two ranks' second-stage sums, identical on both ranks, over
    several calls (both parity slots, every slice's own sequence counter), with no time-out raised."""
    lib = cuda_ops.lib
    world = 2
    nbytes = lib.pctts_bn_sync_bytes(world)
    assert nbytes > 0
    bufs = [torch.zeros((nbytes + 7) // 8, dtype=torch.float64, device='cuda') for _ in range(world)]
    bases = (ctypes.c_ulonglong * world)(*[b.data_ptr() for b in bufs])
    err = torch.zeros(1, dtype=torch.int32, device='cuda')
    streams = [torch.cuda.Stream() for _ in range(world)]
    nblk, count = 7, 1000.0
    g = torch.Generator(device='cuda').manual_seed(3)
    torch.cuda.synchronize()
    for call in range(5):
        parts = [torch.randn(nblk, 2, C, dtype=torch.float64, device='cuda', generator=g) + (10.0 if call % 2 else 0.0) for _ in range(world)]
        for p_ in parts:
            p_[:, 1] = p_[:, 1].abs() * 50 + 100.0          # second moments: keep the variance positive
        outs = []
        torch.cuda.synchronize()
        for r in range(world):
            sums = torch.zeros(2, C, dtype=torch.float64, device='cuda')
            mean, inv = torch.zeros(C, device='cuda'), torch.zeros(C, device='cuda')
            outs.append((sums, mean, inv))
            with torch.cuda.stream(streams[r]):
                rc = lib.pctts_bn_sync_finalize(parts[r].data_ptr(), sums.data_ptr(), nblk, C, count, EPS, MOM, mean.data_ptr(),
                                                inv.data_ptr(), None, None, ctypes.cast(bases, ctypes.c_void_p), r, world,
                                                err.data_ptr(), streams[r].cuda_stream)
                assert rc == 0
        torch.cuda.synchronize()
        assert int(err) == 0
        # expected: per rank the second stage in the kernel's order is a sum over 7 partials (lanes 0..6 of one butterfly) -- compare
        # against the single-replica launch of the same kernel, then add in rank order
        single = []
        for r in range(world):
            s1 = torch.zeros(2, C, dtype=torch.float64, device='cuda')
            m1, i1 = torch.zeros(C, device='cuda'), torch.zeros(C, device='cuda')
            assert lib.pctts_bn_sync_finalize(parts[r].data_ptr(), s1.data_ptr(), nblk, C, count, EPS, MOM, m1.data_ptr(), i1.data_ptr(),
                                              None, None, None, 0, 1, None, torch.cuda.current_stream().cuda_stream) == 0
            single.append(s1)
        torch.cuda.synchronize()
        want = single[0] + single[1]
        for sums, mean, inv in outs:
            assert torch.equal(sums, want)
            mu = want[0] / count
            var = (want[1] / count - mu * mu).clamp_min(0.0)
            eps32 = float(torch.tensor(EPS, dtype=torch.float32))          # the kernel receives eps as a float
            assert torch.equal(mean, mu.float()) and torch.allclose(inv, (1.0 / torch.sqrt(var + eps32)).float(), rtol=2e-7, atol=0)
        assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])
    nsl = (C + 31) // 32
    for b in bufs:
        assert b.view(torch.int64)[:nsl].tolist() == [5] * nsl          # every slice's sequence counter advanced once per call
