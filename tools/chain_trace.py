"""Clock trace of conv2d_chain_kernel (CTA 0): where a layer's time goes.  python tools/chain_trace.py [fwd|dgrad|tangent]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from percivaltts_b200.ops import CudaOps

ops = CudaOps()
mode = sys.argv[1] if len(sys.argv) > 1 else 'fwd'
N, T, F, L = int(os.environ.get('N', 10)), 400, 129, 8
ws = [torch.randn(5, 5, 1 if i == 0 else 4, 4, device='cuda') * 0.2 for i in range(L)]
bs = [torch.randn(4, device='cuda') * 0.1 for _ in range(L)]
x1 = torch.randn(N, T, F, 1, device='cuda')
x4 = torch.randn(N, T, F, 4, device='cuda')
ys = [torch.empty(N, T, F, 4, device='cuda') for _ in range(L)]
acts = [torch.randn(N, T, F, 4, device='cuda') for _ in range(L)]
g = torch.empty(N, T, F, 1, device='cuda')
if mode == 'fwd':
    run = lambda: ops.conv2d_chain(x1, [dict(w=ws[i], bias=bs[i], y=ys[i], flags=2) for i in range(L)])
elif mode == 'dgrad':
    lay = [dict(w=ws[i], flip=True, aux=acts[i - 1], y=ys[i - 1]) for i in range(L - 1, 0, -1)] + [dict(w=ws[0], flip=True, y=g)]
    run = lambda: ops.conv2d_chain(x4, lay)
else:
    run = lambda: ops.conv2d_chain(x1, [dict(w=ws[i], aux=acts[i], y=acts[i]) for i in range(L)])
for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    run()
e1.record()
torch.cuda.synchronize()
print('%s N=%d: %.1f us per launch' % (mode, N, e0.elapsed_time(e1) * 100))
tr = torch.zeros(9 * 16, dtype=torch.int64, device='cuda')
ops.lib.pctts_conv2d_chain_set_trace(tr.data_ptr())
run()
torch.cuda.synchronize()
ops.lib.pctts_conv2d_chain_set_trace(None)
t = tr.cpu().view(9, 16).tolist()
base = t[0][0]
names = ['mma0', 'issued', '-', '-', 'bmat', '-', 'epi', '-', '-', '-', '-', 'hand', 'clus', 'halo']
print('layer ' + ' '.join('%7s' % n for n in names))
for l in range(L):
    print('%5d ' % l + ' '.join('%7d' % ((t[l][k] - base) if t[l][k] else -1) for k in range(14)))
