"""
Host-logic tests (no GPU) of the SURVEY 8(f) rows around the hot path, through the CPU test double of the kernel interface:
  * Generic (FC) model + the plain LSE optimizer (reference optimizertts.py:378-436, modeltts_common.py:36-55);
  * the device-side validation pass of OptimizerTTSWGAN.update_validation_cost (optimizertts_wgan.py:244-268);
  * the generation tail predict -> de-normalise -> exp(f0) / decompress_spectrum / fwbnd2linbnd (modeltts.py:144-205,
    vocoders.py:194-202);
  * the per-batch-shape LRU of the trainer (many distinct T in one training run).
"""
import os

import numpy as np
import pytest
import torch

from oracle import percival_oracle as O
from oracle import pml_oracle as P
from percivaltts_b200 import arch as A
from percivaltts_b200 import backend_b200, data, engine, modeltts_common, networks_critic, optimizertts, optimizertts_wgan, vocoders
from percivaltts_b200.lse import LSEStepper
from percivaltts_b200.percivaltts import configuration
from percivaltts_b200.trainer import DeviceTrainer
from percivaltts_b200.wgan import WGANStepper

from cpu_ops_double import CpuOpsDouble
from test_api_cpu import make_files, tiny_cfg


@pytest.fixture()
def cpu_backend():
    backend_b200.set_ops(CpuOpsDouble(torch.float32))
    yield
    backend_b200.set_ops(None)


def relerr(got, want):
    got, want = torch.as_tensor(got).double(), torch.as_tensor(want).double()
    return float((got - want).abs().max() / (want.abs().max() + 1e-30))


# ---- Generic FC model + LSE ---------------------------------------------------------------------------------------------

def test_generic_fc_lse_step_and_trajectory_match_oracle_fp64():
    dt = torch.float64
    ga = A.GenericArch(20, 9, 5, 6, ['FC', 'FC', 'FC'])
    shapes = O.generic_param_shapes(20, 9, 5, 6, 3)
    assert [(n, tuple(s), t) for n, s, t in shapes] == [(n, tuple(s), t) for n, s, t in A.generic_param_shapes(20, 9, 5, 6, ga.layertypes)]
    w = O.perturb_params(O.init_params(shapes, seed=3))
    ops = CpuOpsDouble(dt)
    net = engine.GenericFC(ops, ga)
    net.params.load(w)
    rng = np.random.RandomState(1)
    X = rng.uniform(-1, 1, size=(3, 17, 20))
    Y = rng.normal(size=(3, 17, 15))
    st = LSEStepper(ops, net)
    st.step(torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt), apply=False)
    r = O.lse_step(None, O.to_torch(w, dt, True), torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt))
    assert abs(float(st.losses[0]) - float(r['loss'])) < 1e-12
    assert relerr(st.buffers(3, 17)['pred'], r['pred']) < 1e-12
    for k, v in r['grads'].items():
        assert relerr(net.params.g[k], v) < 1e-9, k
    for k, v in r['new_stats'].items():
        assert relerr(net.params.p[k], v) < 1e-12, k
    # Adam trajectory (keras defaults of the LSE scheme: lr 10^-3.39794, 0.9, 0.999, eps 1e-8)
    net.params.load(w)
    tr = O.LSETrainer(None, w, dtype=dt)
    for i in range(4):
        Xb = rng.uniform(-1, 1, size=(2, 11 + i, 20))
        Yb = rng.normal(size=(2, 11 + i, 15))
        cost, _ = tr.train_on_batch(Xb, Yb)
        st.step(torch.tensor(Xb, dtype=dt), torch.tensor(Yb, dtype=dt))
        assert abs(float(st.losses[0]) ** 0.5 - cost) < 1e-10
    for k, v in tr.wg.items():
        assert relerr(net.params.p[k], v.detach()) < 1e-8, k


def test_lse_on_the_dcnn_generator_matches_oracle_fp64():
    dt = torch.float64
    kw = dict(ctxsize=20, spec_size=17, nm_size=5, hiddenwidth=8, ctx_nbcnnlayers=1, ctx_winlen=5, gen_nbcnnlayers=3,
              gen_nbfilters=4, gen_winlen=5, spec_freqlen=5)
    ao, a = O.Arch(**kw), A.Arch(**kw)
    w = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    X, Y, _ = O.synthetic_batch(ao, 2, 13, seed=4)
    ops = CpuOpsDouble(dt)
    G = engine.Generator(ops, a)
    G.params.load(w)
    st = LSEStepper(ops, G)
    st.step(torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt), apply=False)
    r = O.lse_step(ao, O.to_torch(w, dt, True), torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt))
    assert abs(float(st.losses[0]) - float(r['loss'])) < 1e-12
    for k, v in r['grads'].items():
        if float(v.abs().max()) > 1e-12:
            assert relerr(G.params.g[k], v) < 1e-9, k


def test_generic_lse_training_loop_through_the_public_api(cpu_backend, tmp_path):
    """The reference's test_run-style LSE training (run.py with OptimizerTTS / errtype LSE): Generic FC model, base-class hooks."""
    lens = [40, 44, 37, 33, 29, 36]
    fids, indir, outdir, wdir = make_files(tmp_path, lens)
    cfg = configuration()
    cfg.arch_hiddenwidth = 4
    cfg.train_batch_size = 2
    cfg.train_min_nbepochs = 1
    cfg.train_max_nbepochs = 2
    cfg.train_cancel_nodecepochs = 2
    cfg.train_batch_lengthmax = 16
    cfg.train_nbepochs_scalewdata = False
    voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
    model = modeltts_common.Generic(425, voc, layertypes=['FC', 'FC', 'FC'], cfgarch=cfg)
    assert model.count_params() == 2195
    opti = optimizertts.OptimizerTTS(cfg, model, errtype='LSE')
    assert abs(opti.cfg.train_lse_learningrate_log10 + 3.39794) < 1e-12 and opti.cfg.train_lse_adam_epsilon_log10 == -8
    fmodel = str(tmp_path / 'generic.h5')
    res = opti.train(indir, outdir, wdir, fids[:4], fids[4:], fmodel)
    assert res['epoch_stopped'] == 2 and np.isfinite(res['best_val'])
    assert os.path.exists(fmodel + '.weights.npz')
    model2 = modeltts_common.Generic(425, voc, fmodel=str(tmp_path / 'generic-last.h5'))      # the last epoch's weights
    x = np.random.RandomState(1).uniform(-1, 1, size=(1, 21, 425)).astype(np.float32)
    assert np.array_equal(model.predict(x), model2.predict(x)) and model2.count_params() == 2195
    opti.cfg.train_max_nbepochs = 3
    res = opti.train(indir, outdir, wdir, fids[:4], fids[4:], fmodel, cont=True)
    assert res['epoch_stopped'] == 3


# ---- validation pass ---------------------------------------------------------------------------------------------------

def test_device_validation_pass_matches_per_utterance_evaluation(cpu_backend):
    cfg = tiny_cfg()
    voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
    model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
    critic = networks_critic.Critic(voc, 425, cfg)
    opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=critic)
    opti.use_graphs = False
    opti.prepare()
    ao = O.tiny_arch()
    rng = np.random.RandomState(3)
    lens = [23, 31, 19]
    X_vals = [rng.uniform(-1, 1, size=(n, 425)).astype(np.float32) for n in lens]
    Y_vals = [rng.normal(size=(n, 83)).astype(np.float32) for n in lens]
    from collections import defaultdict
    costs = defaultdict(list)
    opti.costs_tra_critic_batches = [1.0]
    crit = opti.update_validation_cost(costs, X_vals, Y_vals)
    # reference definitions, per utterance with the oracle (inference-mode BatchNorm)
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
    assert abs(costs['model_rmse_validation'][-1] - (sq / n) ** 0.5) < 1e-4 * (sq / n) ** 0.5
    assert abs(costs['model_validation'][-1] - np.mean(gens)) < 1e-3 * abs(np.mean(gens))
    assert np.isfinite(crit) and costs['critic_validation_ltm'][-1] == crit and costs['critic_training'][-1] == 1.0
    # second epoch reuses the resident validation set
    cache = opti._val_cache
    opti.update_validation_cost(costs, X_vals, Y_vals)
    assert opti._val_cache is cache and len(costs['critic_validation']) == 2


# ---- generation tail -----------------------------------------------------------------------------------------------------

def test_generation_tail_chain_matches_oracle(cpu_backend, tmp_path):
    cfg = tiny_cfg()
    voc = vocoders.VocoderPML(16000, 0.005, 65, 17, dftlen=512)
    model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
    lens = [30, 26]
    fids, indir, outdir, wdir = make_files(tmp_path, lens)
    rng = np.random.RandomState(5)
    mean = rng.normal(size=83).astype(np.float32)
    std = rng.uniform(0.5, 1.5, size=83).astype(np.float32)
    mean.tofile(str(tmp_path / 'cmp' / 'mean4norm.dat'))
    std.tofile(str(tmp_path / 'cmp' / 'std4norm.dat'))
    x = data.load(indir, fids)[0]
    CMP, f0, SPEC, NM = model.generate_features_device(x, mean, std)
    pred = model.predict(x[np.newaxis])[0]
    wantCMP, wf0, wSPEC, wNM = P.generation_tail(pred, mean, std, 16000, 512, 65, 17)
    assert CMP.shape == (31, 83) and f0.shape == (31,) and SPEC.shape == (31, 257) and NM.shape == (31, 257)
    assert np.abs(CMP - wantCMP).max() < 1e-5 * np.abs(wantCMP).max()
    assert np.abs(f0 - wf0).max() < 1e-5 * np.abs(wf0).max()
    assert np.abs(SPEC - wSPEC).max() < 1e-4 * np.abs(wSPEC).max()
    assert np.abs(NM - wNM).max() < 1e-5 * max(np.abs(wNM).max(), 1.0)
    # the loop of modeltts.py:144-205 with an injected synthesiser and writer
    calls, written = [], []

    def synthesizer(fs, f0s, S, N):
        calls.append((fs, f0s.shape, S.shape, N.shape))
        return np.zeros(f0s.shape[0] * 80, dtype=np.float32)

    model.generate_wav(indir, outdir, fids, str(tmp_path / 'syn'), synthesizer=synthesizer,
                       wavwrite=lambda path, wav, fs: written.append(path))
    assert len(calls) == 4 and calls[0][0] == 16000 and calls[1][2] == (30, 257)           # resynthesis + synthesis per utterance
    assert len(written) == 4 and 'F0[Hz]' in voc.features_err and len(voc.features_err['NM']) == 2
    # without a synthesiser: the vocoder inputs go to disk
    model.generate_wav(indir, outdir, fids, str(tmp_path / 'feat'), do_objmeas=False, do_resynth=False)
    got = np.fromfile(str(tmp_path / 'feat' / (fids[1] + '.spec')), dtype=np.float32).reshape(-1, 257)
    assert got.shape[0] == 27 and np.isfinite(got).all()
    with pytest.raises(NotImplementedError):
        voc.synthesis(pred)


# ---- per-shape LRU ----------------------------------------------------------------------------------------------------------

def test_shape_cache_evicts_whole_entries_and_keeps_the_trajectory():
    kw = dict(ctxsize=12, spec_size=9, nm_size=3, hiddenwidth=6, ctx_nbcnnlayers=1, ctx_winlen=3, gen_nbcnnlayers=2,
              gen_nbfilters=2, gen_winlen=3, spec_freqlen=3)
    ao, a = O.Arch(**kw), A.Arch(**kw)
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
    _, wls, wbar = O.wlswgan_weights(ao)
    finals = []
    for max_shapes in (64, 2):
        st = WGANStepper(CpuOpsDouble(torch.float64), a, w_ls=wls, w_wgan_mean=wbar)
        st.G.params.load(wg)
        st.C.params.load(wc)
        tr = DeviceTrainer(st, use_graphs=False, max_shapes=max_shapes)
        for i, T in enumerate([7, 9, 11, 7, 13, 9, 7]):
            X, Y, alpha = O.synthetic_batch(ao, 2, T, seed=20 + i)
            tr.train_on_batch(i * 10, X, Y, alpha=alpha)
        finals.append((st.C.params.flat.clone(), st.G.params.flat.clone(), tr.evictions, len(tr._shapes),
                       len(st.G._plans), len(st.C._plans), len(st._bufs)))
    assert finals[0][2] == 0 and finals[0][3] == 4
    assert finals[1][2] >= 4 and finals[1][3] == 2 and finals[1][4] == 2 and finals[1][6] == 2 and finals[1][5] <= 4
    assert torch.equal(finals[0][0], finals[1][0]) and torch.equal(finals[0][1], finals[1][1])
