"""
BASELINE config #4: critic-only gradient-penalty microbench, B=10, T in {256, 512, 1024, 2048, 4096}.
Timed per T (CUDA events, 3 warm-up + 10 timed, inputs rotate over 3 batches so that the maps do not stay in L2 at large T):
  gp_pass      = interpolation + critic forward on xhat + input gradient + per-sample norm / penalty + tangent sweep +
                 parameter gradients of the penalty (one critic application; optimizertts_wgan.py:44-68)
  critic_step  = the whole critic step (frozen generator, three applications, GP, Adam)
One JSON line per T.
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from percivaltts_b200 import arch as A, wgan as W
from percivaltts_b200.ops import CudaOps
from percivaltts_b200.wgan import WGANStepper


def main():
    ops = CudaOps()
    a = A.Arch()
    st = WGANStepper(ops, a, w_ls=np.ones(a.outsize), w_wgan_mean=0.5)
    st.G.params.load(A.init_params(A.generator_param_shapes(a), 1))
    st.C.params.load(A.init_params(A.critic_param_shapes(a), 2))
    B, S = 10, a.spec_size
    rng = np.random.RandomState(0)
    for T in (256, 512, 1024, 2048, 4096):
        M = B * T
        sets = []
        for i in range(3):
            X = torch.tensor(rng.normal(size=(B, T, a.ctxsize)).astype(np.float32)).cuda()
            Y = torch.tensor(rng.normal(size=(B, T, a.outsize)).astype(np.float32)).cuda()
            F = torch.tensor(rng.normal(size=(B, T, a.outsize)).astype(np.float32)).cuda()
            sets.append((X, Y, F))
        alpha = torch.rand(B, device='cuda')
        cpl = st.C.plan(1, B, T)
        s0 = ops.empty(3 * B, T, S)
        g, ss = ops.empty(B, T, S, 1), ops.zeros(B)
        rows = slice(0, B)

        def gp_pass(X, Y, F):
            with ops.step_scope():
                ops.critic_pack_inputs(Y, F, alpha, s0, S, 1)
                xhat = s0[2 * B:].view(B, T, S, 1)
                d = st.C.forward(cpl, xhat, X)
                ops.fill_segments(cpl['dd'].view(-1), M, 1.0)
                st.C.backward_to_input(cpl, g)
                ops.gp_norms(g.view(B, T * S), ss)
                ops.gp_scale(g.view(B, T * S), ss, xhat.view(B, T * S), st.lam, B, st.losses, W.L_GP)
                st.C.tangent_forward(cpl, rows)
                st.C.param_grads(cpl, first_order_apps=1)

        def timeit(fn, n=10):
            for i in range(3):
                fn(*sets[i % 3])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(n):
                fn(*sets[i % 3])
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / n

        ops.lib.pctts_launch_count.restype = __import__('ctypes').c_long
        n0 = ops.lib.pctts_launch_count()
        gp_pass(*sets[0])
        nl = ops.lib.pctts_launch_count() - n0
        t_gp = timeit(gp_pass)
        t_cs = timeit(lambda X, Y, F: st.critic_step(X, Y))
        print(json.dumps({'config': 'gp_microbench', 'B': B, 'T': T, 'frames': M, 'gp_pass_ms': round(t_gp, 4),
                          'gp_pass_frames_per_s': round(M / t_gp * 1e3), 'gp_pass_launches': int(nl),
                          'critic_step_ms': round(t_cs, 4), 'critic_step_frames_per_s': round(M / t_cs * 1e3),
                          'gp_loss': float(st.losses[W.L_GP]), 'precision': ops.precision, 'eager': True}), flush=True)
        del sets, s0, g
        st.C._plans.clear(); st.G._plans.clear(); st._bufs.clear()
        ops._shadows.clear(); ops._persist.clear(); ops._red.clear(); ops._wg.clear()
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
