"""Developer probe: device time of the tcgen05 GEMM for a few shapes / operand majors / precisions (CUDA events, warm L2)."""
import os, sys, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from percivaltts_b200.ops import CudaOps

def timeit(fn, iters=20):
    """10 launches captured in a CUDA graph (no host launch overhead in the number), replayed `iters` times."""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (10 * iters) * 1e3   # us

shapes = [(4096, 256, 8192), (4096, 256, 256)] if os.environ.get('PCTTS_TC_DEBUG') else [(4096, 256, 8192), (4096, 256, 256), (12000, 256, 516), (4096, 1024, 256)]
for prec in ('bf16', 'fp32x3', 'fp32'):
    ops = CudaOps(precision=prec)
    for (M, N, K), (ta, tb) in itertools.product(shapes, [(False, True)]):
        A = torch.randn((K, M) if ta else (M, K), device='cuda')
        B = torch.randn((N, K) if tb else (K, N), device='cuda')
        C = torch.empty(M, N, device='cuda')
        with ops.step_scope():
            ops.gemm(A, B, C, transA=ta, transB=tb)          # splits happen here; shadows stay fresh inside the scope
            us = timeit(lambda: ops.gemm(A, B, C, transA=ta, transB=tb))
        fl = 2.0 * M * N * K
        print('%-7s M=%5d N=%4d K=%5d tA=%d tB=%d  %8.1f us  %7.1f TFLOP/s(fp32-equiv) x%d products' %
              (prec, M, N, K, ta, tb, us, fl / us / 1e6, ops.nterms), flush=True)
