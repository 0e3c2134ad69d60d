"""Developer probe: which operands are split into bf16 planes during one critic step and one generator step (demo shape)."""
import os, sys, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from percivaltts_b200 import arch as A
from percivaltts_b200.ops import CudaOps
from percivaltts_b200.wgan import WGANStepper

ops = CudaOps()
a = A.Arch()
st = WGANStepper(ops, a, errtype='WGAN')
B, T = 10, 400
X = torch.randn(B, T, a.ctxsize, device='cuda')
Y = torch.randn(B, T, a.outsize, device='cuda')
for _ in range(2):
    st.critic_step(X, Y)
    st.generator_step(X, Y)
log = []
orig = ops._split
def traced(t, sh):
    log.append((ops._lane, tuple(t.shape), 'persist' if ops._is_persistent(t.data_ptr()) else 'act'))
    return orig(t, sh)
ops._split = traced
for name, fn in (('critic', st.critic_step), ('generator', st.generator_step)):
    del log[:]
    fn(X, Y)
    torch.cuda.synchronize()
    c = collections.Counter(log)
    print('%s step: %d splits' % (name, len(log)))
    for k, n in sorted(c.items(), key=lambda kv: (kv[0][2], kv[0][0], -kv[0][1][0] * kv[0][1][-1])):
        print('   lane %d %-8s %-22s x%d' % (k[0], k[2], k[1], n))
