"""Developer probe: per-tensor deviation of the GPU training trajectory from the fp64 oracle (tests/test_step_gpu.py set-up)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch
from oracle import percival_oracle as O
from percivaltts_b200 import wgan as W
from percivaltts_b200.ops import CudaOps
import test_step_gpu as TS

ops = CudaOps(precision=os.environ.get('PREC', 'fp32'))
ao, wg, wc, wgt, wct, X, Y, alpha, st = TS.setup(ops, TS.MID, 2, 40)
tr = O.OracleTrainer(ao, wg, wc, dtype=torch.float64)
gu = 0
nb = int(os.environ.get('NB', 6))
for batchid in range(nb):
    X, Y, al = O.synthetic_batch(ao, 2, 40, seed=100 + batchid)
    bid = batchid * 5
    tr.train_on_batch(bid, X, Y, al)
    Xd, Yd = torch.tensor(X).cuda(), torch.tensor(Y).cuda()
    st.critic_step(Xd, Yd, alpha=torch.tensor(al).cuda())
    if bid % W.critic_runs(gu) == 0:
        st.generator_step(Xd, Yd)
        gu += 1
    for name, store, ref, lr in (('C', st.C.params, tr.wc, st.copt[0]), ('G', st.G.params, tr.wg, st.gopt[0])):
        num = den = 0.0
        worst = []
        for k, v in ref.items():
            got, want = store.p[k].detach().cpu().double(), v.detach().double()
            n_, d_ = float(((got - want) ** 2).sum()), float((want ** 2).sum())
            num += n_; den += d_
            worst.append((n_, k, float((got - want).abs().max()) / lr, want.numel()))
        worst.sort(reverse=True)
        print('batch %d %s relL2 %.3e   top: %s' % (batchid, name, (num / den) ** 0.5,
              ', '.join('%s sq=%.2e max=%.2f lr n=%d' % (k, n_, m, c) for n_, k, m, c in worst[:4])))
