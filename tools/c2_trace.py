"""Developer probe: clock trace of CTA 0 of the tcgen05 Conv2D kernel at the batched critic map (30 x 400 x 129 x 4)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from percivaltts_b200.ops import CudaOps

ops = CudaOps(precision=os.environ.get('PREC', 'fp32'))
nb = int(os.environ.get('NB', 30))
x, w = torch.randn(nb, 400, 129, 4, device='cuda'), torch.randn(5, 5, 4, 4, device='cuda')
y = torch.empty_like(x)
for _ in range(2):
    ops.conv2d(x, w, y)
tr = torch.zeros(32 * 8, dtype=torch.int64, device='cuda')
ops.lib.pctts_conv2d_tc_set_trace(tr.data_ptr())
flush = torch.empty(64 * 1024 * 1024, device='cuda')
if os.environ.get('FLUSH'):
    flush.zero_()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ops.conv2d(x, w, y)
e1.record()
torch.cuda.synchronize()
print('event time %.2f us' % (1e3 * e0.elapsed_time(e1)))
ops.lib.pctts_conv2d_tc_set_trace(None)
t = tr.cpu().view(32, 8)
t0 = int(t[0, 0])
print('tile: prod_start stage_free filled | mma_ready mma_issued | epi_start acc_ready stored   (cycles since first stamp)')
for i in range(32):
    if int(t[i, 0]) == 0:
        break
    print(i, [int(v) - t0 if int(v) else None for v in t[i]])

print('kernel entry -> roles start: %d cycles; entry -> exit: %d cycles' % (int(t[31, 1] - t[31, 0]), int(t[31, 2] - t[31, 0])))
print('prologue stamps rel. entry: setup %d, producers start %d, weight operand built %d' % tuple(int(t[31, k] - t[31, 0]) for k in (3, 4, 1)))
print('entry rel. first stamp', int(t[31, 0]) - t0)
