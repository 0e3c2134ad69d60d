"""
GPU parity of the SURVEY 8(f) rows around the hot path (all kernels through the C ABI) against the CPU oracle:
Generic FC model + LSE step, the device-side validation pass, the generation tail chained on the device, and the
bounded per-batch-shape cache of the trainer.
"""
from collections import defaultdict

import numpy as np
import pytest
import torch

from oracle import percival_oracle as O
from oracle import pml_oracle as P
from percivaltts_b200 import arch as A
from percivaltts_b200 import backend_b200, engine, modeltts_common, networks_critic, optimizertts_wgan, vocoders
from percivaltts_b200.lse import LSEStepper
from percivaltts_b200.trainer import DeviceTrainer
from percivaltts_b200.wgan import WGANStepper
from test_api_cpu import tiny_cfg

pytestmark = pytest.mark.gpu

TOL = 1e-3
MID = dict(ctxsize=60, spec_size=33, nm_size=9, hiddenwidth=32, ctx_nbcnnlayers=1, ctx_winlen=21, gen_nbcnnlayers=4,
           gen_nbfilters=4, gen_winlen=5, spec_freqlen=5)


def relerr(got, want):
    got = torch.as_tensor(got).detach().cpu().double()
    want = torch.as_tensor(want).detach().cpu().double()
    return float((got - want).abs().max() / (want.abs().max() + 1e-30))


@pytest.mark.parametrize('ctx,S,NM,H,L,B,T', [(425, 65, 17, 4, 3, 2, 293), (425, 129, 33, 256, 3, 4, 200), (60, 33, 9, 32, 1, 3, 57)])
def test_generic_fc_lse_step_vs_oracle(cuda_ops, ctx, S, NM, H, L, B, T):
    """Generic(['FC']*L) + LSE (optimizertts.py:386-402; modeltts_common.py:36-55): loss, prediction, gradients, BN moving
    statistics and four Adam updates against the fp64 oracle."""
    dt = torch.float64
    ga = A.GenericArch(ctx, S, NM, H, ['FC'] * L)
    w = O.perturb_params(O.init_params(O.generic_param_shapes(ctx, S, NM, H, L), seed=3))
    net = engine.GenericFC(cuda_ops, ga)
    net.params.load(w)
    rng = np.random.RandomState(1)
    X = rng.uniform(-1, 1, size=(B, T, ctx)).astype(np.float32)
    Y = rng.normal(size=(B, T, 1 + S + NM)).astype(np.float32)
    st = LSEStepper(cuda_ops, net)
    st.step(torch.tensor(X).cuda(), torch.tensor(Y).cuda(), apply=False)
    torch.cuda.synchronize()
    r = O.lse_step(None, O.to_torch(w, dt, True), torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt))
    assert abs(float(st.losses[0]) - float(r['loss'])) < TOL * float(r['loss'])
    assert relerr(st.buffers(B, T)['pred'], r['pred']) < TOL
    num = den = 0.0
    for k, v in r['grads'].items():
        got = net.params.g[k].detach().cpu().double()
        num += float(((got - v) ** 2).sum())
        den += float((v ** 2).sum())
        assert relerr(got, v) < 5e-2, k            # per tensor: LeakyReLU mask flips allowed (tests/test_step_gpu.py: check_grads)
    assert (num / den) ** 0.5 < TOL
    for k, v in r['new_stats'].items():
        assert relerr(net.params.p[k], v) < TOL, k
    net.params.load(w)
    tr, tr32 = O.LSETrainer(None, w, dtype=dt), O.LSETrainer(None, w, dtype=torch.float32)
    for i in range(4):
        Xb = rng.uniform(-1, 1, size=(B, T, ctx)).astype(np.float32)
        Yb = rng.normal(size=(B, T, 1 + S + NM)).astype(np.float32)
        cost, _ = tr.train_on_batch(Xb, Yb)
        tr32.train_on_batch(Xb, Yb)
        st.step(torch.tensor(Xb).cuda(), torch.tensor(Yb).cuda())
        assert abs(float(st.losses[0]) ** 0.5 - cost) < TOL * cost
    # Four Adam updates move every weight by about lr each, whatever its gradient: an element whose gradient sits at round-off
    # level steps the other way in ANY fp32 evaluation.  The yardstick is therefore the oracle's own fp32 run of the same four
    # batches: the distance to the fp64 trajectory (L2 over the trained weights) within 3x that run's (or 1e-3 of the weights'
    # norm), and the number of elements more than half an Adam step away within 3x its count (+ 0.1 % of the elements).
    lr = st.opt[0]
    d_gpu = d_32 = nrm = 0.0
    off_gpu = off_32 = n = 0
    for k, v in tr.wg.items():
        if not v.requires_grad:
            continue
        v = v.detach().double()
        e_gpu = (net.params.p[k].detach().cpu().double() - v).abs()
        e_32 = (tr32.wg[k].detach().double() - v).abs()
        d_gpu += float((e_gpu ** 2).sum())
        d_32 += float((e_32 ** 2).sum())
        nrm += float((v ** 2).sum())
        off_gpu += int((e_gpu > 0.5 * lr).sum())
        off_32 += int((e_32 > 0.5 * lr).sum())
        n += v.numel()
    assert d_gpu ** 0.5 <= max(3.0 * d_32 ** 0.5, TOL * nrm ** 0.5), (d_gpu ** 0.5, d_32 ** 0.5, nrm ** 0.5)
    assert off_gpu <= 3 * off_32 + max(1, n // 1000), (off_gpu, off_32, n)


def test_generation_tail_on_device_vs_oracle(cuda_ops):
    """predict -> de-normalise (mean4norm / std4norm, modeltts.py:154-165) -> exp(f0) / decompress_spectrum / NM fwbnd2linbnd
    (vocoders.py:194-202) chained on the device with one D2H: PML tolerance 1e-5 against the numpy oracle on the same prediction."""
    backend_b200.set_ops(cuda_ops)
    try:
        cfg = tiny_cfg()
        voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
        model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
        rng = np.random.RandomState(5)
        mean = rng.normal(size=83).astype(np.float32)
        std = rng.uniform(0.5, 1.5, size=83).astype(np.float32)
        for T in (1, 3, 4, 293):                                            # TMA spans of 4 frames + tails
            x = rng.uniform(-1, 1, size=(T, 425)).astype(np.float32)
            n0 = cuda_ops.lib.pctts_launch_count()
            CMP, f0, SPEC, NM = model.generate_features_device(x, mean, std)
            assert cuda_ops.lib.pctts_launch_count() > n0
            pred = model.predict(x[np.newaxis])[0]
            wCMP, wf0, wSPEC, wNM = P.generation_tail(pred, mean, std, 16000, 4096, 65, 17)
            assert SPEC.shape == (T, 2049) and NM.shape == (T, 2049) and f0.shape == (T,) and CMP.shape == (T, 83)
            assert np.abs(CMP - wCMP).max() <= 1e-5 * np.abs(wCMP).max()
            assert np.abs(f0 / wf0 - 1).max() <= 1e-5
            assert np.abs(SPEC / wSPEC - 1).max() <= 1e-5 * max(1.0, np.abs(np.log(wSPEC)).max())   # 1e-5 on the log-amplitude
            assert np.abs(NM - wNM).max() <= 1e-5 * max(1.0, np.abs(wNM).max())
            # the host-side de-normalisation + two decompress calls (round 1) agree to fp32 round-off (device fma vs numpy mul + add)
            f0b, SPECb, NMb = voc.decompress_features(pred * std + mean)
            assert np.allclose(SPECb, SPEC, rtol=2e-5, atol=0) and np.allclose(NMb, NM, rtol=1e-5, atol=1e-6)
    finally:
        backend_b200.set_ops(None)


def test_device_validation_pass_vs_oracle(cuda_ops):
    """update_validation_cost (optimizertts_wgan.py:244-268) as one device-resident pass with a single D2H: prediction RMSE and
    generator objective against the oracle's inference-mode evaluation, utterance by utterance."""
    backend_b200.set_ops(cuda_ops)
    try:
        cfg = tiny_cfg()
        voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
        model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
        critic = networks_critic.Critic(voc, 425, cfg)
        opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=critic)
        opti.prepare()
        ao = O.tiny_arch()
        rng = np.random.RandomState(3)
        lens = [293, 453, 311, 293]
        X_vals = [rng.uniform(-1, 1, size=(n, 425)).astype(np.float32) for n in lens]
        Y_vals = [rng.normal(size=(n, 83)).astype(np.float32) for n in lens]
        costs = defaultdict(list)
        opti.costs_tra_critic_batches = [1.0]
        crit = opti.update_validation_cost(costs, X_vals, Y_vals)
        wg = O.to_torch(model.kerasmodel.get_weights(), torch.float64)
        wc = O.to_torch(critic.kerasmodel.get_weights(), torch.float64)
        _, wls, wbar = O.wlswgan_weights(ao)
        sq, n, gens = 0.0, 0, []
        for X, Y in zip(X_vals, Y_vals):
            Xt, Yt = torch.tensor(X[None], dtype=torch.float64), torch.tensor(Y[None], dtype=torch.float64)
            with torch.no_grad():
                pred = O.generator_forward(ao, wg, Xt, training=False)
                d = O.critic_forward(ao, wc, pred, Xt)
            sq += float(((Yt - pred) ** 2).sum())
            n += Yt.numel()
            gens.append(wbar * float((-d).mean()) + float((((Yt - pred) ** 2) * torch.tensor(wls)).mean()))
        assert abs(costs['model_rmse_validation'][-1] - (sq / n) ** 0.5) < TOL * (sq / n) ** 0.5
        assert abs(costs['model_validation'][-1] - np.mean(gens)) < TOL * abs(np.mean(gens))
        assert np.isfinite(crit)
        h0 = opti.trainer.d2h_bytes
        opti.update_validation_cost(costs, X_vals, Y_vals)               # resident set, same numbers but for the random alphas
        assert abs(costs['model_rmse_validation'][-1] - costs['model_rmse_validation'][-2]) < 1e-6
        assert abs(costs['model_validation'][-1] - costs['model_validation'][-2]) < 1e-6 * abs(costs['model_validation'][-1])
    finally:
        backend_b200.set_ops(None)


def test_shape_cache_bounds_device_memory_over_many_batch_lengths(cuda_ops):
    """A training run meets many T (shortest utterance of each batch): the per-shape state (plans, step buffers, operand shadows,
    reduction workspaces, CUDA graphs) is one LRU entry and is freed together, so device memory stays bounded."""
    ao, a = O.Arch(**MID), A.Arch(**MID)
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
    _, wls, wbar = O.wlswgan_weights(ao)
    st = WGANStepper(cuda_ops, a, w_ls=wls, w_wgan_mean=wbar)
    st.G.params.load(wg)
    st.C.params.load(wc)
    tr = DeviceTrainer(st, max_shapes=4)
    mem = []
    for i in range(32):
        T = 64 + 8 * (i % 16) if i < 16 else 64 + 8 * ((5 * i) % 16)      # 16 distinct lengths, then revisits (graphs get captured)
        X, Y, _ = O.synthetic_batch(ao, 2, T, seed=i)
        c, g = tr.train_on_batch(i * 5, X, Y)
        assert np.isfinite(c)
        torch.cuda.synchronize()
        mem.append(torch.cuda.memory_allocated())
    assert len(tr._shapes) <= 4 and tr.evictions >= 12
    assert len(st.G._plans) <= 4 and len(st._bufs) <= 4 and len(st.C._plans) <= 8
    # bounded: the footprint after 32 batches over 16 lengths stays within 1.5x of what the first 6 batches reached
    assert max(mem[6:]) <= 1.5 * max(mem[:6]), (max(mem[:6]), max(mem[6:]))
