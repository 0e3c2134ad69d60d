"""Developer timing of the eager (non-graph) critic / generator steps at the demo size, CUDA events."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from percivaltts_b200 import arch as A
from percivaltts_b200.ops import CudaOps
from percivaltts_b200.wgan import WGANStepper

def main():
    B, T = int(os.environ.get('B', 10)), int(os.environ.get('T', 400))
    ops = CudaOps()
    a = A.Arch()
    st = WGANStepper(ops, a, w_ls=np.ones(a.outsize), w_wgan_mean=0.5)
    st.G.params.load(A.init_params(A.generator_param_shapes(a), 1))
    st.C.params.load(A.init_params(A.critic_param_shapes(a), 2))
    rng = np.random.RandomState(0)
    X = torch.tensor(rng.normal(size=(B, T, a.ctxsize)).astype(np.float32)).cuda()
    Y = torch.tensor(rng.normal(size=(B, T, a.outsize)).astype(np.float32)).cuda()
    def timeit(fn, n=5):
        for _ in range(2): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    ops.launches = 0
    st.critic_step(X, Y); nc = ops.launches
    ops.launches = 0
    st.generator_step(X, Y); ng = ops.launches
    print('critic step   %.3f ms  (%d launches)' % (timeit(lambda: st.critic_step(X, Y)), nc))
    print('generator step %.3f ms  (%d launches)' % (timeit(lambda: st.generator_step(X, Y)), ng))
    gpl = st.G.plan(B, T)
    out = st.buffers(B, T)['fake']
    print('G forward     %.3f ms' % timeit(lambda: st.G.forward(gpl, X, out, training=True)))
    print('losses', st.losses.cpu().numpy())
    # per-phase profile with torch profiler-free event timing of individual kernels of interest
    H = a.hiddenwidth
    zx = gpl['zx']; U = [st.G.params.p['g.f0.blstm.fw.recurrent'], st.G.params.p['g.f0.blstm.bw.recurrent']]
    print('lstm fwd      %.3f ms' % timeit(lambda: ops.lstm_fwd(zx, U, gpl['hout'], gpl['gates'], gpl['cstate'])))
    print('lstm bwd      %.3f ms' % timeit(lambda: ops.lstm_bwd(gpl['dhout'], U, gpl['gates'], gpl['cstate'], gpl['dzl'])))
    from percivaltts_b200.engine import _im2col_view
    xp = gpl['ctx_xp'][0]; k = st.G.params.p['g.ctx.cnn0.kernel']
    t = timeit(lambda: ops.gemm(_im2col_view(xp, T, a.ctx_winlen), k.view(-1, H), gpl['ctx_z'][0]))
    print('conv1d gemm   %.3f ms  %.1f TFLOP/s' % (t, 2.0 * B * T * H * k.numel() / H / t / 1e9))
    t = timeit(lambda: ops.gemm(_im2col_view(xp, T, a.ctx_winlen), gpl['ctx_z'][0], st.G.params.g['g.ctx.cnn0.kernel'].view(-1, H), transA=True))
    print('conv1d wgrad  %.3f ms' % t)
    cpl = st.C.plan(3, B, T)
    t = timeit(lambda: ops.conv2d(cpl['a'][0], st.C.params.p['c.spec.cnn1.kernel'], cpl['a'][1], bias=st.C.params.p['c.spec.cnn1.bias'], flags=2))
    print('conv2d 4->4 (3B) %.3f ms' % t)
    t = timeit(lambda: ops.conv2d_wgrad(cpl['a'][0], cpl['da'][1], st.C.params.g['c.spec.cnn1.kernel']))
    print('conv2d wgrad (3B) %.3f ms' % t)

if __name__ == '__main__':
    main()
