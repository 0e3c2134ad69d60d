"""Full-size (B=10, T=400) critic step: GP input gradient and critic gradients against the fp64 oracle, with the chained Conv2D
stack on / off, beside the fp32 oracle's own distance to fp64 (mask-flip yardstick)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import percival_oracle as O
from percivaltts_b200 import arch as A, wgan as W
from percivaltts_b200.ops import CudaOps
from percivaltts_b200.wgan import WGANStepper

def rel(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30)), float(((a - b) ** 2).sum().sqrt() / (b ** 2).sum().sqrt())

B, T = int(os.environ.get('B', 10)), int(os.environ.get('T', 400))
kw = A.Arch().as_dict()
ao, a = O.Arch(**kw), A.Arch(**kw)
wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
X, Y, alpha = O.synthetic_batch(ao, B, T, seed=5)
_, wls, wbar = O.wlswgan_weights(ao)
dt = torch.float64
rc = O.critic_step(ao, O.to_torch(wg, dt, True), O.to_torch(wc, dt, True), torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt), torch.tensor(alpha, dtype=dt).reshape(-1, 1, 1))
rc32 = O.critic_step(ao, O.to_torch(wg, torch.float32, True), O.to_torch(wc, torch.float32, True), torch.tensor(X), torch.tensor(Y), torch.tensor(alpha).reshape(-1, 1, 1))
S = ao.spec_size
print('fp32 oracle vs fp64: gp_grad max-rel %.2e L2 %.2e' % rel(rc32['gp_grad'][:, :, 1:1 + S], rc['gp_grad'][:, :, 1:1 + S]))
for chain in (True, False):
    ops = CudaOps()
    ops.conv_chain = chain
    st = WGANStepper(ops, a, w_ls=wls, w_wgan_mean=wbar)
    st.G.params.load(wg); st.C.params.load(wc)
    st.critic_step(torch.tensor(X).cuda(), torch.tensor(Y).cuda(), alpha=torch.tensor(alpha).cuda(), apply=False)
    torch.cuda.synchronize()
    print('chain=%d: gp_grad max-rel %.2e L2 %.2e; loss %.6f (oracle %.6f)' % ((int(chain),) + rel(st.buffers(B, T)['g'].view(B, T, -1), rc['gp_grad'][:, :, 1:1 + S]) + (float(st.losses[W.L_CRITIC]), float(rc['loss']))))
    worst = [(k,) + rel(st.C.params.g[k], v) + rel(rc32['grads'][k], v) for k, v in rc['grads'].items() if float(v.abs().max()) > 1e-10]
    worst.sort(key=lambda r: -r[1])
    for r in worst[:6]:
        print('   %-24s gpu max-rel %.2e L2 %.2e | fp32 oracle max-rel %.2e L2 %.2e' % r)
