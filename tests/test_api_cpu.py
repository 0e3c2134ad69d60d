"""
Host-logic tests of the reference-facing API (no GPU): the python-3 mirror of percivaltts' model / critic / optimizer /
vocoder / data interfaces, driven through the CPU test double of the kernel interface.  Reads like the reference's
tests/test_smoke.py and tests/test_smoke_tensorflowkeras.py.
"""
import os

import numpy as np
import pytest
import torch

from oracle import percival_oracle as O
from percivaltts_b200 import backend_b200, data, modeltts_common, networks_critic, optimizertts_wgan, vocoders
from percivaltts_b200.percivaltts import configuration, is_int

from cpu_ops_double import CpuOpsDouble


def tiny_cfg():
    cfg = configuration()             # tests/test_smoke_tensorflowkeras.py:183-190
    cfg.arch_hiddenwidth = 2
    cfg.arch_ctx_nbcnnlayers = 2
    cfg.arch_ctx_winlen = 3
    cfg.arch_gen_nbcnnlayers = 2
    cfg.arch_gen_nbfilters = 2
    cfg.arch_gen_winlen = 3
    cfg.arch_spec_freqlen = 3
    cfg.train_batch_size = 2
    cfg.fs = 16000
    return cfg


@pytest.fixture()
def cpu_backend():
    backend_b200.set_ops(CpuOpsDouble(torch.float32))
    yield
    backend_b200.set_ops(None)


def test_utils_and_sizes():
    assert is_int('12') and not is_int('a1')                      # tests/test_smoke.py:39-40
    v = vocoders.VocoderPML(16000, 0.005, 65, 17)
    assert v.featuressize() == 83 and v.specsize() == 65 and v.noisesize() == 17 and v.vuvsize() == -1 and v.f0size() == 1
    assert vocoders.VocoderWORLD(16000, 0.005, 65, 17).featuressize() == 84             # test_smoke_tensorflowkeras.py:149
    wins = [[-0.5, 0.0, 0.5], [1.0, -2.0, 1.0]]
    assert vocoders.VocoderPML(16000, 0.005, 65, 17, mlpg_wins=wins).featuressize() == 249    # :133
    assert vocoders.VocoderWORLD(16000, 0.005, 65, 17, mlpg_wins=wins).featuressize() == 252  # :158


def test_count_params_known_answer():
    cfg = configuration()
    cfg.arch_hiddenwidth = 4
    model = modeltts_common.Generic(425, vocoders.VocoderPML(16000, 0.005, 65, 17), layertypes=['FC', 'FC', 'FC'], cfgarch=cfg)
    assert model.count_params() == 2195                           # tests/test_smoke_tensorflowkeras.py:51-53
    with pytest.raises(ValueError):
        modeltts_common.Generic(425, vocoders.VocoderPML(16000, 0.005, 65, 17), layertypes=['FC', 'XX'], cfgarch=cfg)


def test_getpathandshape_cases(tmp_path):
    assert data.getpathandshape('dummy.fwlspec') == ('dummy.fwlspec', None)                  # tests/test_smoke.py:61-72
    assert data.getpathandshape('dummy.fwlspec:(-1,129)') == ('dummy.fwlspec', (-1, 129))
    assert data.getpathandshape('dummy.fwlspec:(-1,129)', (-1, 12)) == ('dummy.fwlspec', (-1, 12))
    assert data.getpathandshape('dummy.fwlspec', (-1, 12)) == ('dummy.fwlspec', (-1, 12))
    assert data.getlastdim('dummy.fwlspec') == 1 and data.getlastdim('dummy.fwlspec:(-1,129)') == 129
    np.zeros(7, dtype=np.float32).tofile(str(tmp_path / 'mean.dat'))
    assert data.getpathandshape(str(tmp_path / '*.x') + ':(-1,mean.dat)')[1] == (-1, 7)


def make_files(tmp_path, lens, ctx=425, out=83):
    rng = np.random.RandomState(0)
    fids = ['utt%02d' % i for i in range(len(lens))]
    for d in ('lab', 'cmp', 'w'):
        os.makedirs(str(tmp_path / d), exist_ok=True)
    for fid, n in zip(fids, lens):
        rng.uniform(-1, 1, size=(n + 1, ctx)).astype(np.float32).tofile(str(tmp_path / 'lab' / (fid + '.lab')))
        Y = rng.normal(size=(n, out)).astype(np.float32)
        Y[:, 66:] = rng.uniform(0, 1, size=(n, out - 66))
        Y.tofile(str(tmp_path / 'cmp' / (fid + '.cmp')))
        w = np.ones((n, 1), dtype=np.float32)
        w[:n // 10] = 0
        w[-(n // 10):] = 0
        w.tofile(str(tmp_path / 'w' / (fid + '.w')))
    return (fids, str(tmp_path / 'lab' / '*.lab') + ':(-1,%d)' % ctx, str(tmp_path / 'cmp' / '*.cmp') + ':(-1,%d)' % out,
            str(tmp_path / 'w' / '*.w') + ':(-1,1)')


def test_data_layer(tmp_path):
    lens = [66, 74, 63, 49, 29, 58, 59, 45, 61, 59]
    fids, indir, outdir, wdir = make_files(tmp_path, lens)
    Xs = data.load(indir, fids)
    assert len(Xs) == 10 and Xs[0].shape == (67, 425) and data.gettotallen(Xs) == sum(lens) + 10
    Ys, Ws = data.load(outdir, fids), data.load(wdir, fids)
    Xs, Ys, Ws = data.croplen([Xs, Ys, Ws])
    assert all(x.shape[0] == y.shape[0] == n for x, y, n in zip(Xs, Ys, lens))
    [Xs, Ys], Ws = data.croplen_weight([Xs, Ys], Ws, thresh=0.5)
    assert all(float(w.min()) > 0.5 for w in Ws)
    for mode in ('begend', 'begendbigger', 'all'):
        np.random.seed(1)
        X, Y, W = data.load_inoutset(indir, outdir, wdir, fids, length=None, lengthmax=20, maskpadtype='randshift', cropmode=mode)
        assert X.shape == (10, 20, 425) and Y.shape == (10, 20, 83) and W.shape == (10, 20, 1) and X.dtype == np.float32
    X, Y, W = data.load_inoutset(indir, outdir, wdir, fids[:3], maskpadtype='padright')
    assert X.shape[1] == max(n - 2 * (n // 10) for n in lens[:3]) - 1
    assert data.cost_0pred_rmse(Ys) > 0


def test_wgan_optimizer_hot_path_matches_oracle(cpu_backend):
    """prepare() + train_on_batch() through the public API = the oracle's train_on_batch trajectory."""
    cfg = tiny_cfg()
    voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
    model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
    critic = networks_critic.Critic(voc, 425, cfg)
    assert model.count_params() == O.count_params(O.generator_param_shapes(O.tiny_arch()))
    opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=critic)
    assert opti.cfg.train_wgan_pg_lambda == 10 and opti.cfg.train_batch_size == 2 and opti.cfg.train_wgan_LScoef == 0.25
    opti.prepare()
    ao = O.tiny_arch()
    w_wgan, w_ls, wbar = opti.wganls_weights()
    ow, ols, owbar = O.wlswgan_weights(ao)
    assert np.allclose(w_wgan, ow) and np.allclose(w_ls, ols) and abs(wbar - owbar) < 1e-15
    tr = O.OracleTrainer(ao, model.kerasmodel.get_weights(), critic.kerasmodel.get_weights(), dtype=torch.float64)
    for batchid in range(3):
        X, Y, alpha = O.synthetic_batch(ao, 2, 31 + batchid, seed=50 + batchid)
        bid = 10 * batchid
        oc, og, _, _ = tr.train_on_batch(bid, X, Y, alpha)
        cost = opti.train_on_batch(bid, X, Y, alpha=alpha)
        assert abs(opti.costs_tra_critic_batches[-1] - oc) < 2e-3 * max(abs(oc), 1e-2)
        assert (cost is None) == (og is None)
        if og is not None:
            assert abs(cost - og) < 2e-3 * abs(og)
    assert opti.generator_updates == 3
    assert opti.train_on_batch(7, X, Y) is None                  # 7 % 10 != 0: critic-only batch
    got = model.kerasmodel.get_weights()
    for k, v in tr.wg.items():
        assert np.abs(got[k] - v.detach().numpy()).max() < 5e-3 * (np.abs(v.detach().numpy()).max() + 1e-6), k
    # predict: inference mode, one utterance at a time (modeltts.py:68-69)
    x = X[:1]
    want = O.generator_forward(ao, O.to_torch(got, torch.float64), torch.tensor(x, dtype=torch.float64), training=False)
    assert np.abs(model.predict(x) - want.numpy()).max() < 1e-4


def test_full_training_loop_checkpoint_and_resume(cpu_backend, tmp_path):
    """The reference's tiny WLSWGAN smoke run (tests/test_smoke_tensorflowkeras.py:183-203): a few epochs of the base
    loop on 8 training + 2 validation utterances, model files written, `cont=True` resumes."""
    lens = [40, 44, 37, 33, 29, 36, 41, 30, 35, 38]
    fids, indir, outdir, wdir = make_files(tmp_path, lens)
    cfg = tiny_cfg()
    cfg.train_min_nbepochs = 1
    cfg.train_max_nbepochs = 2
    cfg.train_cancel_nodecepochs = 2
    cfg.train_batch_lengthmax = 16
    cfg.train_nbepochs_scalewdata = False
    voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
    model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
    critic = networks_critic.Critic(voc, 425, cfg)
    opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=critic)
    opti.use_graphs = False
    fmodel = str(tmp_path / 'model.h5')
    res = opti.train(indir, outdir, wdir, fids[:8], fids[8:], fmodel)
    assert res['epoch_stopped'] == 2 and np.isfinite(res['best_val'])
    for suffix in ('.arch.json', '.weights.npz', '.cfgextras.pkl'):
        assert os.path.exists(fmodel + suffix)
    # save -> load round trip keeps cfg and extras (tests/test_smoke_tensorflowkeras.py:77-79)
    model2 = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
    cfg_loaded, extras = model2.load(fmodel)
    assert 'cost_val' in extras
    w1, w2 = model.kerasmodel.get_weights(), model2.kerasmodel.get_weights()
    assert all(np.array_equal(w1[k], w2[k]) for k in w1)
    x = np.random.RandomState(1).uniform(-1, 1, size=(1, 21, 425)).astype(np.float32)
    assert np.array_equal(model.predict(x), model2.predict(x))
    # resume for one more epoch
    opti.cfg.train_max_nbepochs = 3
    res = opti.train(indir, outdir, wdir, fids[:8], fids[8:], fmodel, cont=True)
    assert res['epoch_stopped'] == 3
    model.generate_cmp(indir, str(tmp_path / 'out' / '*.cmp'), fids[8:])
    gen = np.fromfile(str(tmp_path / 'out' / (fids[8] + '.cmp')), dtype=np.float32).reshape(-1, 83)
    assert gen.shape[0] == lens[8] + 1 and np.isfinite(gen).all()


def test_nan_cost_raises(cpu_backend, tmp_path):
    lens = [30, 31, 32, 33]
    fids, indir, outdir, wdir = make_files(tmp_path, lens)
    cfg = tiny_cfg()
    cfg.train_min_nbepochs = 1
    cfg.train_max_nbepochs = 1
    voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
    model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
    w = model.kerasmodel.get_weights()
    w['g.spec.out.bias'][:] = np.nan
    model.kerasmodel.set_weights(w)
    opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=networks_critic.Critic(voc, 425, cfg))
    with pytest.raises(ValueError):
        opti.train(indir, outdir, wdir, fids[:2], fids[2:], str(tmp_path / 'm.h5'))
