"""
Step-level parity of the reduced-precision GEMM modes (BASELINE config #3 arithmetic) against the fp64 oracle.

north_star states the bf16 tolerance separately from the fp32 one; DESIGN.md section 2 fixes it as
    losses / outputs within 2e-2 relative, cosine >= 0.995 between the flat gradient and the oracle's
    (measured on B200, demo architecture at B=4, T=96: losses 1e-3 .. 3e-3, critic gradient cosine 0.99990, generator gradient
    cosine 0.9976 -- the generator's BatchNorm backward at a small batch differences large terms, which is also why the default
    mode carries 24 significand bits)
for `bf16` (one bf16 plane, one tcgen05 product per k-slice, fp32 accumulation in TMEM, fp32 master weights, fp32 BatchNorm
statistics and LSTM state), and for `fp32x3` (two planes / three products) 1e-3 on losses and outputs with gradient cosine
>= 0.99999 (the per-tensor 1e-3 of the default six-product mode is NOT claimed for it: DESIGN.md section 4.7).
"""
import pytest
import torch

from oracle import percival_oracle as O
from percivaltts_b200 import arch as A
from percivaltts_b200 import wgan as W
from percivaltts_b200.wgan import WGANStepper

pytestmark = pytest.mark.gpu

MID = dict(ctxsize=60, spec_size=33, nm_size=9, hiddenwidth=32, ctx_nbcnnlayers=1, ctx_winlen=21, gen_nbcnnlayers=4,
           gen_nbfilters=4, gen_winlen=5, spec_freqlen=5)
DEMO = A.Arch().as_dict()

# precision -> (loss / output tolerance, minimum cosine of the flat gradients)
TOLERANCES = {'bf16': (2e-2, 0.995), 'fp32x3': (1e-3, 0.99999)}


def relerr(got, want):
    got = torch.as_tensor(got).detach().cpu().double()
    want = torch.as_tensor(want).detach().cpu().double()
    return float((got - want).abs().max() / (want.abs().max() + 1e-30))


def flat_cosine(store, want):
    num = a2 = b2 = 0.0
    for k, v in want.items():
        got = store.g[k].detach().cpu().double()
        v = v.detach().double()
        num += float((got * v).sum())
        a2 += float((got * got).sum())
        b2 += float((v * v).sum())
    return num / ((a2 * b2) ** 0.5 + 1e-300)


@pytest.mark.parametrize('precision', ['bf16', 'fp32x3'])
@pytest.mark.parametrize('kw,B,T', [(MID, 3, 57), (DEMO, 4, 96)])
def test_reduced_precision_step_vs_oracle(precision, kw, B, T):
    if not torch.cuda.is_available():
        pytest.skip('no CUDA device')
    from percivaltts_b200.ops import CudaOps
    ops = CudaOps(precision=precision)
    tol, mincos = TOLERANCES[precision]
    ao, a = O.Arch(**kw), A.Arch(**kw)
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
    X, Y, alpha = O.synthetic_batch(ao, B, T, seed=5)
    _, wls, wbar = O.wlswgan_weights(ao)
    st = WGANStepper(ops, a, w_ls=wls, w_wgan_mean=wbar)
    st.G.params.load(wg)
    st.C.params.load(wc)
    dt = torch.float64
    wgt, wct = O.to_torch(wg, dt, True), O.to_torch(wc, dt, True)
    Xt, Yt, al = torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt), torch.tensor(alpha, dtype=dt)
    rc = O.critic_step(ao, wgt, wct, Xt, Yt, al.reshape(-1, 1, 1))
    rg = O.generator_step(ao, wgt, wct, Xt, Yt)
    Xd, Yd = torch.tensor(X).cuda(), torch.tensor(Y).cuda()
    st.critic_step(Xd, Yd, alpha=torch.tensor(alpha).cuda(), apply=False)
    torch.cuda.synchronize()
    losses = st.losses.cpu()
    report = {}
    for slot, key in ((W.L_REAL, 'l_real'), (W.L_FAKE, 'l_fake'), (W.L_GP, 'gp'), (W.L_CRITIC, 'loss')):
        report[key] = abs(float(losses[slot]) - float(rc[key])) / max(abs(float(rc[key])), 1e-2)
    report['critic_cos'] = flat_cosine(st.C.params, rc['grads'])
    st.generator_step(Xd, Yd, apply=False)
    torch.cuda.synchronize()
    losses = st.losses.cpu()
    report['pred'] = relerr(st.buffers(B, T)['fake'], rg['pred'])
    report['gen_loss'] = abs(float(losses[W.L_GEN]) - float(rg['loss'])) / abs(float(rg['loss']))
    report['gen_cos'] = flat_cosine(st.G.params, rg['grads'])
    print('precision %s B=%d T=%d: %s' % (precision, B, T, ' '.join('%s=%.3g' % kv for kv in sorted(report.items()))))
    for key in ('l_real', 'l_fake', 'gp', 'loss', 'pred', 'gen_loss'):
        assert report[key] < tol, (key, report)
    assert report['critic_cos'] >= mincos and report['gen_cos'] >= mincos, report
