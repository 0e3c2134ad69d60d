"""
Host-logic tests (no GPU): the hand-derived forward / backward / double-backward schedules of
percivaltts_b200.engine and .wgan, executed through the CPU test double of the kernel interface
(tests/cpu_ops_double.py), must reproduce the oracle's autograd results.

fp64 run: agreement to ~1e-12 proves the derivation (GP double backprop, BN backward, BPTT, shared ctx branch);
fp32 run: the tolerance of BASELINE.json's north_star (1e-3 relative) with margin.
"""
import numpy as np
import pytest
import torch

from oracle import percival_oracle as O
from percivaltts_b200 import arch as A
from percivaltts_b200 import wgan as W
from percivaltts_b200.wgan import WGANStepper

from cpu_ops_double import CpuOpsDouble

SMALL = dict(ctxsize=20, spec_size=17, nm_size=5, hiddenwidth=8, ctx_nbcnnlayers=1, ctx_winlen=5, gen_nbcnnlayers=3,
             gen_nbfilters=4, gen_winlen=5, spec_freqlen=5)
DEEPCTX = dict(ctxsize=12, spec_size=9, nm_size=3, hiddenwidth=6, ctx_nbcnnlayers=3, ctx_winlen=4, gen_nbcnnlayers=2,
               gen_nbfilters=2, gen_winlen=3, spec_freqlen=3)
NOCTXCNN = dict(ctxsize=12, spec_size=9, nm_size=3, hiddenwidth=6, ctx_nbcnnlayers=0, ctx_winlen=3, gen_nbcnnlayers=1,
                gen_nbfilters=2, gen_winlen=3, spec_freqlen=3)


def make_case(arch_kwargs, B, T, dtype, errtype='WLSWGAN'):
    ao = O.Arch(**arch_kwargs)
    a = A.Arch(**arch_kwargs)
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
    X, Y, alpha = O.synthetic_batch(ao, B, T, seed=5)
    wgt = O.to_torch(wg, dtype, requires_grad=True)
    wct = O.to_torch(wc, dtype, requires_grad=True)
    Xt, Yt = torch.tensor(X, dtype=dtype), torch.tensor(Y, dtype=dtype)
    _, wls, wbar = O.wlswgan_weights(ao)
    st = WGANStepper(CpuOpsDouble(dtype), a, errtype=errtype, w_ls=wls, w_wgan_mean=wbar)
    st.G.params.load(wg)
    st.C.params.load(wc)
    return ao, wgt, wct, Xt, Yt, torch.tensor(alpha, dtype=dtype), st


def relerr(got, want):
    return float((got - want).abs().max() / (want.abs().max() + 1e-30))


def check_grads(store, want, tol):
    for k, v in want.items():
        if float(v.abs().max()) < 1e-12:
            assert float(store.g[k].abs().max()) < 1e-9, k
        else:
            assert relerr(store.g[k], v) < tol, (k, relerr(store.g[k], v))


@pytest.mark.parametrize('arch_kwargs,B,T', [
    (A.Arch(**{k: getattr(O.tiny_arch(), k) for k in A.Arch.FIELDS}).as_dict(), 2, 37),
    (SMALL, 3, 23), (DEEPCTX, 2, 11), (NOCTXCNN, 2, 7), (SMALL, 1, 1)])
def test_steps_match_oracle_fp64(arch_kwargs, B, T):
    ao, wgt, wct, Xt, Yt, alpha, st = make_case(arch_kwargs, B, T, torch.float64)
    rc = O.critic_step(ao, wgt, wct, Xt, Yt, alpha.reshape(-1, 1, 1))
    st.critic_step(Xt.clone(), Yt.clone(), alpha=alpha, apply=False)
    for slot, key in ((W.L_REAL, 'l_real'), (W.L_FAKE, 'l_fake'), (W.L_GP, 'gp'), (W.L_CRITIC, 'loss')):
        assert abs(float(st.losses[slot]) - float(rc[key])) < 1e-10 * max(1.0, abs(float(rc[key])))
    check_grads(st.C.params, rc['grads'], 1e-9)
    rg = O.generator_step(ao, wgt, wct, Xt, Yt)
    st.generator_step(Xt.clone(), Yt.clone(), apply=False)
    assert abs(float(st.losses[W.L_GEN]) - float(rg['loss'])) < 1e-10
    assert relerr(st.buffers(B, T)['fake'], rg['pred']) < 1e-10
    check_grads(st.G.params, rg['grads'], 1e-9)
    for k, v in rg['new_stats'].items():
        assert relerr(st.G.params.p[k], v) < 1e-10, k


def test_steps_match_oracle_fp32_tolerance():
    """north_star tolerance: outputs, losses and gradients within 1e-3 relative in fp32."""
    ao, wgt, wct, Xt, Yt, alpha, st = make_case(SMALL, 4, 50, torch.float32)
    rc = O.critic_step(ao, wgt, wct, Xt, Yt, alpha.reshape(-1, 1, 1))
    st.critic_step(Xt.clone(), Yt.clone(), alpha=alpha, apply=False)
    assert abs(float(st.losses[W.L_CRITIC]) - float(rc['loss'])) < 1e-3 * abs(float(rc['loss']))
    check_grads(st.C.params, rc['grads'], 1e-3)
    rg = O.generator_step(ao, wgt, wct, Xt, Yt)
    st.generator_step(Xt.clone(), Yt.clone(), apply=False)
    assert abs(float(st.losses[W.L_GEN]) - float(rg['loss'])) < 1e-3 * abs(float(rg['loss']))
    check_grads(st.G.params, rg['grads'], 1e-3)


def test_wgan_errtype_and_training_trajectory():
    """Five batches of train_on_batch policy (critic every batch, generator when batchid % critic_runs == 0) with Adam:
    parameters after the trajectory match the oracle trainer (optimizertts_wgan.py:216-241)."""
    for errtype in ('WGAN', 'WLSWGAN'):
        ao, wgt, wct, Xt, Yt, alpha, st = make_case(SMALL, 2, 19, torch.float64, errtype=errtype)
        wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
        wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
        tr = O.OracleTrainer(ao, wg, wc, errtype=errtype, dtype=torch.float64)
        gen_updates = 0
        for batchid in range(3):
            X, Y, al = O.synthetic_batch(ao, 2, 19, seed=100 + batchid)
            tr.train_on_batch(batchid * 10, X, Y, al)       # batchid*10 -> generator trained on every call
            Xd, Yd = torch.tensor(X, dtype=torch.float64), torch.tensor(Y, dtype=torch.float64)
            st.critic_step(Xd, Yd, alpha=torch.tensor(al, dtype=torch.float64))
            if (batchid * 10) % W.critic_runs(gen_updates) == 0:
                st.generator_step(Xd, Yd)
                gen_updates += 1
        for k, v in tr.wc.items():
            assert relerr(st.C.params.p[k], v.detach()) < 1e-7, k
        for k, v in tr.wg.items():
            assert relerr(st.G.params.p[k], v.detach()) < 1e-7, k


def test_validation_losses_match_oracle_inference_mode():
    ao, wgt, wct, Xt, Yt, alpha, st = make_case(SMALL, 1, 31, torch.float64)
    with torch.no_grad():
        fake = O.generator_forward(ao, wgt, Xt, training=False)
    rc = O.critic_step(ao, wgt, wct, Xt, Yt, alpha.reshape(-1, 1, 1), gen_training=False)
    st.evaluate_losses(Xt.clone(), Yt.clone(), alpha=alpha)
    assert abs(float(st.losses[W.L_CRITIC]) - float(rc['loss'])) < 1e-10
    _, wls, wbar = O.wlswgan_weights(ao)
    d_fake = O.critic_forward(ao, wct, fake, Xt)
    want = wbar * float((-d_fake).mean()) + float((((Yt - fake) ** 2) * torch.tensor(wls)).mean())
    assert abs(float(st.losses[W.L_GEN]) - want) < 1e-10
    pred = st.predict(Xt.clone())
    assert relerr(pred, fake) < 1e-12


def test_critic_runs_schedule():
    assert [W.critic_runs(i) for i in (0, 24, 25, 499, 500, 501, 1000)] == [10, 10, 5, 5, 10, 5, 10]


@pytest.mark.parametrize('B,T', [(3, 23), (1, 1)])
def test_steps_with_the_chained_conv_stack_match_oracle_fp64(B, T):
    """The critic's Conv2D stack issued as multi-layer chains (ops.conv2d_chain: forward, data gradient with on-the-fly kernel
    flip and the input gradient as last layer, in-place tangent sweep) -- same results as the per-layer schedule."""
    ao, wgt, wct, Xt, Yt, alpha, st = make_case(SMALL, B, T, torch.float64)
    st.ops.conv_chain = True
    assert st.C._chain(T)
    rc = O.critic_step(ao, wgt, wct, Xt, Yt, alpha.reshape(-1, 1, 1))
    st.critic_step(Xt.clone(), Yt.clone(), alpha=alpha, apply=False)
    for slot, key in ((W.L_REAL, 'l_real'), (W.L_FAKE, 'l_fake'), (W.L_GP, 'gp'), (W.L_CRITIC, 'loss')):
        assert abs(float(st.losses[slot]) - float(rc[key])) < 1e-10 * max(1.0, abs(float(rc[key])))
    check_grads(st.C.params, rc['grads'], 1e-9)
    rg = O.generator_step(ao, wgt, wct, Xt, Yt)
    st.generator_step(Xt.clone(), Yt.clone(), apply=False)
    assert abs(float(st.losses[W.L_GEN]) - float(rg['loss'])) < 1e-10
    check_grads(st.G.params, rg['grads'], 1e-9)
    ao, wgt, wct, Xt, Yt, alpha, st = make_case(SMALL, B, T, torch.float64)       # fresh BatchNorm moving statistics
    st.ops.conv_chain = True
    st.evaluate_losses(Xt.clone(), Yt.clone(), alpha=alpha)
    rv = O.critic_step(ao, wgt, wct, Xt, Yt, alpha.reshape(-1, 1, 1), gen_training=False)
    assert abs(float(st.losses[W.L_CRITIC]) - float(rv['loss'])) < 1e-10 * max(1.0, abs(float(rv['loss'])))
