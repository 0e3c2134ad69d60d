import os, sys, torch
sys.path.insert(0, '/root/repo')
from percivaltts_b200.ops import CudaOps
o = CudaOps()
M, H = 12000, 256
p2 = torch.randn(M, H, device='cuda'); w = torch.randn(H, 1, device='cuda'); d = torch.empty(M, 1, device='cuda')
dd = torch.randn(M, 1, device='cuda'); dp = torch.empty(M, H, device='cuda'); gw = torch.empty(H, 1, device='cuda'); b = torch.zeros(1, device='cuda')
flush = torch.empty(64 * 1024 * 1024, device='cuda')
def t(fn, name):
    for _ in range(3): fn()
    tot = 0
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize(); tot += e0.elapsed_time(e1)
    print('%-10s %.1f us' % (name, 100 * tot))
t(lambda: o.gemm(p2, w, d, bias=b), 'gemv')
t(lambda: o.gemm(dd, w, dp, transB=True, aux=p2), 'outer')
t(lambda: o.gemm(p2, dd, gw, transA=True), 'wcolsum')
