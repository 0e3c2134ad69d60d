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


def run(ops=None, Ts=(256, 512, 1024, 2048, 4096), iters=10, graphs=True):
    """-> list of per-T dicts.  graphs=True: the GP pass of each of the three rotating input sets is captured in a CUDA graph
    (as the training step is) and the replays are timed; the whole critic step is timed eagerly beside it."""
    import ctypes
    ops = ops or CudaOps()
    a = A.Arch()
    st = WGANStepper(ops, a, w_ls=np.ones(a.outsize), w_wgan_mean=0.5)
    st.G.params.load(A.init_params(A.generator_param_shapes(a), 1))
    st.C.params.load(A.init_params(A.critic_param_shapes(a), 2))
    B, S = 10, a.spec_size
    rng = np.random.RandomState(0)
    ops.lib.pctts_launch_count.restype = ctypes.c_long
    results = []
    for T in Ts:
        M = B * T
        sets = []
        for i in range(3):
            X = torch.tensor(rng.normal(size=(B, T, a.ctxsize)).astype(np.float32)).cuda()
            Y = torch.tensor(rng.normal(size=(B, T, a.outsize)).astype(np.float32)).cuda()
            F = torch.tensor(rng.normal(size=(B, T, a.outsize)).astype(np.float32)).cuda()
            sets.append((X, Y, F))
        alpha = torch.rand(B, device='cuda')
        tag = ('gp', B, T)
        with ops.shape_scope(tag):
            cpl = st.C.plan(1, B, T)
            s0 = ops.empty(3 * B, T, S)
            g, ss = ops.empty(B, T, S, 1), ops.zeros(B)
            rows = slice(0, B)

            def gp_pass(X, Y, F):
                with ops.shape_scope(tag), ops.step_scope():
                    ops.critic_pack_inputs(Y, F, alpha, s0, S, 1)
                    xhat = s0[2 * B:].view(B, T, S, 1)
                    st.C.forward(cpl, xhat, X)
                    ops.fill_segments(cpl['dd'].view(-1), M, 1.0)
                    st.C.backward_to_input(cpl, g)
                    ops.gp_norms(g.view(B, T * S), ss)
                    ops.gp_scale(g.view(B, T * S), ss, xhat.view(B, T * S), st.lam, B, st.losses, W.L_GP)
                    st.C.tangent_forward(cpl, rows)
                    st.C.param_grads(cpl, first_order_apps=1)

            def timeit(fn, n=iters):
                for i in range(3):
                    fn(i % 3)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(n):
                    fn(i % 3)
                e1.record()
                torch.cuda.synchronize()
                return e0.elapsed_time(e1) / n

            n0 = ops.lib.pctts_launch_count()
            gp_pass(*sets[0])
            nl = ops.lib.pctts_launch_count() - n0
            torch.cuda.synchronize()
            if graphs:
                gr = []
                for i in range(3):
                    gi = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gi):
                        gp_pass(*sets[i])
                    gr.append(gi)
                t_gp = timeit(lambda i: gr[i].replay())
            else:
                t_gp = timeit(lambda i: gp_pass(*sets[i]))
            gp_loss = float(st.losses[W.L_GP])
            t_cs = timeit(lambda i: st.critic_step(sets[i][0], sets[i][1]), n=max(3, iters // 2))
        results.append({'config': 'gp_microbench', 'B': B, 'T': T, 'frames': M, 'gp_pass_ms': round(t_gp, 4),
                        'gp_pass_frames_per_s': round(M / t_gp * 1e3), 'gp_pass_launches': int(nl),
                        'critic_step_ms_eager': round(t_cs, 4), 'critic_step_frames_per_s': round(M / t_cs * 1e3),
                        'gp_loss': gp_loss, 'precision': ops.precision, 'graph': bool(graphs)})
        if graphs:
            del gr
        del sets, s0, g, cpl
        st.C._plans.clear()
        st.release(B, T)
        ops.release_shape(tag)
        torch.cuda.empty_cache()
    return results


def main():
    for r in run():
        print(json.dumps(r), flush=True)


if __name__ == '__main__':
    main()
