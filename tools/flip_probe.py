"""Developer probe: per-tensor relative error of the critic gradients (GPU fp32 vs fp64 oracle) over several seeds."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import percival_oracle as O
from percivaltts_b200 import arch as A
from percivaltts_b200.ops import CudaOps
from percivaltts_b200.wgan import WGANStepper
ops = CudaOps()
kw = A.Arch().as_dict()
ao, a = O.Arch(**kw), A.Arch(**kw)
for seed, B, T in ((5, 2, 48), (6, 2, 48), (7, 2, 40), (8, 3, 32)):
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
    X, Y, alpha = O.synthetic_batch(ao, B, T, seed=seed)
    _, wls, wbar = O.wlswgan_weights(ao)
    st = WGANStepper(ops, a, w_ls=wls, w_wgan_mean=wbar)
    st.G.params.load(wg); st.C.params.load(wc)
    dt = torch.float64
    rc = O.critic_step(ao, O.to_torch(wg, dt, True), O.to_torch(wc, dt, True), torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt),
                       torch.tensor(alpha, dtype=dt).reshape(-1, 1, 1))
    st.critic_step(torch.tensor(X).cuda(), torch.tensor(Y).cuda(), alpha=torch.tensor(alpha).cuda(), apply=False)
    errs = []
    for k, v in rc['grads'].items():
        if float(v.abs().max()) < 1e-10: continue
        g = st.C.params.g[k].cpu().double()
        errs.append((float((g - v).abs().max() / v.abs().max()), k, float(v.abs().max())))
    errs.sort(reverse=True)
    print(seed, B, T, ['%s %.2e (|v| %.1e)' % (k, e, m) for e, k, m in errs[:4]])
