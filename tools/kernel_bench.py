"""
Per-kernel timing at the demo shapes (B=10, T=400): every hot kernel of libpctts.so alone, CUDA events on the launching
stream, L2 flushed (1 GiB fill, which also hides the host enqueue time) before every timed launch, against its roofline (MEASURED_PEAKS.json).
One JSON line per kernel on stdout:  {"kernel", "ms", "achieved", "unit", "peak", "frac", "work"}.

    python tools/kernel_bench.py [--only substring] [--iters 20]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from percivaltts_b200 import arch as A
from percivaltts_b200 import sigproc as sp
from percivaltts_b200.ops import CudaOps, EPI_LRELU

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--only', default='')
    ap.add_argument('--iters', type=int, default=20)
    ap.add_argument('--precision', default=None)
    args = ap.parse_args()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except (OSError, ValueError):
        pass
    hbm, tf = float(peaks.get('hbm_gbs', 6650.0)), float(peaks.get('bf16_tflops', 1590.0))
    ops = CudaOps(precision=args.precision)
    simt = CudaOps(precision='simt')
    dev = ops.device
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.float32, device=dev)   # 1 GiB: also long enough to hide the host-side enqueue
    g = torch.Generator(device='cuda').manual_seed(1)

    def rnd(*shape):
        return torch.randn(*shape, device=dev, generator=g)

    def timeit(fn):
        for _ in range(3):
            fn()
        tot = 0.0
        for _ in range(args.iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / args.iters

    def report(name, fn, nbytes=None, flops=None, note=''):
        if args.only and args.only not in name:
            return
        ms = timeit(fn)
        rec = {'kernel': name, 'ms': round(ms, 5)}
        if nbytes is not None:
            rec.update(bound='hbm', achieved=round(nbytes / ms / 1e6, 1), unit='GB/s', peak=hbm, frac=round(nbytes / ms / 1e6 / hbm, 4),
                       work='%.3f MB algorithmic' % (nbytes / 1e6))
        if flops is not None:
            rec.update(bound='tensor', achieved=round(flops / ms / 1e9, 2), unit='TFLOP/s', peak=tf, frac=round(flops / ms / 1e9 / tf, 4),
                       work='%.3f GFLOP algorithmic' % (flops / 1e9))
        if note:
            rec['note'] = note
        print(json.dumps(rec), flush=True)

    B, T, S, H, W, CTX, O = 10, 400, 129, 256, 21, 425, 163
    M = B * T
    # ---- Conv2D 4->4 5x5 on the batched critic map (3B) and on the generator map (B): read + write of the map
    for nb, tag in ((3 * B, '3B'), (B, 'B')):
        x, y, w, bias = rnd(nb, T, S, 4), torch.empty(nb, T, S, 4, device=dev), rnd(5, 5, 4, 4) * 0.2, rnd(4)
        mapb = nb * T * S * 4 * 4
        fl = 2.0 * nb * T * S * 100 * 4
        report('conv2d_tc_fwd 4->4 %s (%s)' % (tag, ops.precision), lambda: ops.conv2d(x, w, y, bias=bias, flags=EPI_LRELU), nbytes=2 * mapb,
               note='%.2f GFLOP' % (fl / 1e9))
        report('conv2d_simt_fwd 4->4 %s' % tag, lambda: simt.conv2d(x, w, y, bias=bias, flags=EPI_LRELU), nbytes=2 * mapb)
        dw = torch.empty(5, 5, 4, 4, device=dev)
        dbias = torch.empty(4, device=dev)
        report("conv2d_wgrad 4->4 %s (+bias)" % tag, lambda: ops.conv2d_wgrad(x, y, dw, dbias=dbias), nbytes=2 * mapb, note="%.2f GFLOP, CUDA cores" % (fl / 1e9))
    # ---- weight + bias gradients of the critic's whole stack (1->4 and 7 x 4->4) in one launch: bound by the fp32 FMA pipe
    #      (148 SMs x 128 FMA/clk x 1.965 GHz = 37.2 T FMA/s), 400 FMA per pixel and layer
    for nb, tag in ((2 * B, '2B'), (B, 'B')):
        xs = [rnd(nb, T, S, 1)] + [rnd(nb, T, S, 4) for _ in range(7)]
        dys = [rnd(nb, T, S, 4) for _ in range(8)]
        dws = [torch.empty(5, 5, xx.shape[3], 4, device=dev) for xx in xs]
        dbs = [torch.empty(4, device=dev) for _ in xs]
        lay = [dict(x=xs[i], dy=dys[i], dw=dws[i], dbias=dbs[i]) for i in range(8)]
        if (not args.only or 'wgrad_multi' in args.only) and ops.conv2d_wgrad_multi_ok(lay):
            ms = timeit(lambda: ops.conv2d_wgrad_multi(lay))
            fma = 8.0 * nb * T * S * 400
            print(json.dumps({'kernel': 'conv2d_wgrad_multi 8 layers %s' % tag, 'ms': round(ms, 5), 'bound': 'fma', 'achieved': round(fma / ms / 1e9, 2),
                              'unit': 'TFMA/s', 'peak': 37.2, 'frac': round(fma / ms / 1e9 / 37.2, 4),
                              'work': '%.3f GFMA issued (the 1->4 layer runs as 4->4 with three zero channels)' % (fma / 1e9)}), flush=True)
    x1, y4, w14 = rnd(3 * B, T, S, 1), torch.empty(3 * B, T, S, 4, device=dev), rnd(5, 5, 1, 4)
    report('conv2d_simt_fwd 1->4 3B', lambda: ops.conv2d(x1, w14, y4), nbytes=3 * B * T * S * 4 * 5)
    w41 = rnd(5, 5, 4, 1)
    report('conv2d_simt_fwd 4->1 3B', lambda: ops.conv2d(y4, w41, x1), nbytes=3 * B * T * S * 4 * 5)
    # ---- Conv1D k=21 425->256: forward and weight gradient (tensor)
    xp, k, z = rnd(B, T + W - 1, CTX), rnd(W, CTX, H) * 0.05, torch.empty(M, H, device=dev)
    with ops.step_scope():
        report('conv1d_tc_fwd', lambda: ops.conv1d(xp, k, z, T), flops=2.0 * M * W * CTX * H)
        dk = torch.empty(W, CTX, H, device=dev)
        report('conv1d_tc_wgrad', lambda: ops.conv1d_wgrad(xp, z, dk, T), flops=2.0 * M * W * CTX * H)
        a, wfc, c = rnd(M, H), rnd(H, H) * 0.05, torch.empty(M, H, device=dev)
        report('dense 256x256 M=4000', lambda: ops.gemm(a, wfc, c), flops=2.0 * M * H * H)
        a3, w772, c3 = rnd(3 * M, 516), rnd(516, H) * 0.05, torch.empty(3 * M, H, device=dev)
        report('dense 516x256 M=12000', lambda: ops.gemm(a3, w772, c3), flops=2.0 * 3 * M * 516 * H)
        wl, zl = rnd(H, 4 * H) * 0.05, torch.empty(M, 4 * H, device=dev)
        report('lstm input projection 256x1024', lambda: ops.gemm(a, wl, zl), flops=2.0 * M * H * 4 * H)
    report('split_bf16 [4000,256]', lambda: ops._split(a, ops.shadow(a)), nbytes=M * H * (4 + 2 * 3))
    # ---- BLSTM recurrence
    zx = [rnd(B, T, 4 * H) * 0.5 for _ in range(2)]
    U = [rnd(H, 4 * H) * 0.05 for _ in range(2)]
    hout = torch.empty(B, T, 2 * H, device=dev)
    gates = [torch.empty(B, T, 4 * H, device=dev) for _ in range(2)]
    cst = [torch.empty(B, T, H, device=dev) for _ in range(2)]
    dz = [torch.empty(B, T, 4 * H, device=dev) for _ in range(2)]
    dh = rnd(B, T, 2 * H)
    if not args.only or 'lstm' in args.only:
        ms = timeit(lambda: ops.lstm_fwd(zx, U, hout, gates, cst))
        print(json.dumps({'kernel': 'lstm_fwd B=10 T=400 H=256 x2 dirs', 'ms': round(ms, 4), 'bound': 'latency', 'us_per_step': round(1e3 * ms / T, 3)}), flush=True)
        ms = timeit(lambda: ops.lstm_bwd(dh, U, gates, cst, dz))
        print(json.dumps({'kernel': 'lstm_bwd B=10 T=400 H=256 x2 dirs', 'ms': round(ms, 4), 'bound': 'latency', 'us_per_step': round(1e3 * ms / T, 3)}), flush=True)
    # ---- BatchNorm / reductions / elementwise (HBM)
    zmap, ymap = rnd(M * S, 4), torch.empty(M * S, 4, device=dev)
    report('bn_stats [516000,4]', lambda: ops.bn_stats(zmap), nbytes=M * S * 16)
    mean, istd, gam, bet = torch.zeros(4, device=dev), torch.ones(4, device=dev), torch.ones(4, device=dev), torch.zeros(4, device=dev)
    report('bn_apply [516000,4]', lambda: ops.bn_apply(zmap, ymap, mean, istd, gam, bet, 1e-3, 0, 1), nbytes=2 * M * S * 16)
    n = 3781084
    p_, g_, m_, v_ = rnd(n), rnd(n), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    step = torch.zeros(1, dtype=torch.int32, device=dev)
    report('adam generator (3.78 M)', lambda: ops.adam(p_, g_, m_, v_, 1e-3, 0.5, 0.9, 1e-7, step), nbytes=7 * 4 * n)
    real, fake, alpha, s0 = rnd(B, T, O), rnd(B, T, O), torch.rand(B, device=dev), torch.empty(3 * B, T, S, device=dev)
    report('critic_pack_inputs', lambda: ops.critic_pack_inputs(real, fake, alpha, s0, S, 1), nbytes=M * S * 4 * 5)
    gsp, ss = rnd(B, T * S), torch.zeros(B, device=dev)
    losses = torch.zeros(8, device=dev)
    report('gp_norms', lambda: ops.gp_norms(gsp, ss), nbytes=M * S * 4)
    r0 = torch.empty(B, T * S, device=dev)
    report('gp_scale', lambda: ops.gp_scale(gsp, ss, r0, 10.0, B, losses, 2), nbytes=2 * M * S * 4)
    wls, dpred = torch.ones(O, device=dev), torch.empty(B, T, O, device=dev)
    report('lse_loss_grad', lambda: ops.lse_loss_grad(fake, real, wls, dpred, 1.0, 1.0, losses, 5), nbytes=3 * M * O * 4)
    # ---- PML band compression (BASELINE config #5, on a 100k-frame chunk: 0.82 GB in)
    if not args.only or 'pml' in args.only:
        TF, NB = 100000, 2049
        spec = torch.rand(TF, NB, device=dev) + 0.01
        for K, log_abs in ((129, 1), (33, 0)):
            idx, wgt = sp.compress_table(16000, 4096, K)
            idx_d, wgt_d = torch.tensor(idx, dtype=torch.int32, device=dev), torch.tensor(wgt, dtype=torch.float32, device=dev)
            out = torch.empty(TF, K, device=dev)
            report('pml_linbnd2fwbnd K=%d' % K, lambda: ops.pml_linbnd2fwbnd(spec, idx_d, wgt_d, out, log_abs), nbytes=TF * (NB + K) * 4)
            idx2, wgt2 = sp.decompress_table(16000, 4096, K)
            i2, w2 = torch.tensor(idx2, dtype=torch.int32, device=dev), torch.tensor(wgt2, dtype=torch.float32, device=dev)
            back = torch.empty(TF, NB, device=dev)
            report('pml_fwbnd2linbnd K=%d' % K, lambda: ops.pml_fwbnd2linbnd(out, i2, w2, back, int(log_abs), int(log_abs)), nbytes=TF * (NB + K) * 4)


if __name__ == '__main__':
    main()
