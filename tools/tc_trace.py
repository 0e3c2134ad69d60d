"""Developer probe: per-k-block clock stamps of the producer / MMA warps of CTA (0,0,0) of one tcgen05 GEMM launch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from percivaltts_b200.ops import CudaOps
prec = sys.argv[1] if len(sys.argv) > 1 else 'fp32'
M, N, K = 4096, 256, int(sys.argv[2]) if len(sys.argv) > 2 else 2048
ops = CudaOps(precision=prec)
A, B, C = torch.randn(M, K, device='cuda'), torch.randn(N, K, device='cuda'), torch.empty(M, N, device='cuda')
buf = torch.zeros(64 * 8, dtype=torch.int64, device='cuda')
with ops.step_scope():
    for _ in range(3): ops.gemm(A, B, C, transB=True)
    torch.cuda.synchronize()
    ops.lib.pctts_gemm_tc_set_trace(buf.data_ptr())
    ops.gemm(A, B, C, transB=True)
    torch.cuda.synchronize()
    ops.lib.pctts_gemm_tc_set_trace(None)
t = buf.cpu().view(64, 8)
t0 = int(t[0, 0])
print(prec, 'k-block: producer[wait_empty_start, got_empty, issued]  mma[wait_full_start, got_full, issued+committed]  (cycles since start)')
print('entry %d  setup_done %d  drained %d  parked %d  barrier %d  stored %d  dealloc %d' % tuple(int(x) - t0 for x in (t[0, 3], t[0, 7], t[1, 3], t[2, 7], t[3, 3], t[1, 7], t[2, 3])))
for i in range(min(K // 64, 8)):
    r = [int(x) - t0 if int(x) else -1 for x in t[i]]
    print('%3d  P %7d %7d %7d   M %7d %7d %7d' % (i, r[0], r[1], r[2], r[4], r[5], r[6]))
