"""
The input pipeline either side of train_on_batch: the per-rank shard loader returns exactly the rank's slice of the
whole batch with exactly the same draws from numpy's global RNG, and the epoch loop's loader thread leaves the order of those
draws -- hence every batch -- as the reference's batch-by-batch loop (optimizertts.py:231-232) produces them.
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, 'golden'))

from percivaltts_b200 import data


@pytest.mark.parametrize('world', [1, 2, 3, 6])
def test_shard_loader_equals_the_slice_of_the_whole_batch(tmp_path, world):
    import data_scenarios as S
    fids = S.make_files(str(tmp_path))
    indir = os.path.join(str(tmp_path), 'lab', '*.lab') + ':(-1,%d)' % S.CTX
    outdir = os.path.join(str(tmp_path), 'cmp', '*.cmp') + ':(-1,%d)' % S.OUT
    wdir = os.path.join(str(tmp_path), 'w', '*.w') + ':(-1,1)'
    state = np.random.get_state()
    try:
        cases = [dict(), dict(lengthmax=20, maskpadtype='randshift'), dict(length=16, maskpadtype='randshift', cropmode='begendbigger'),
                 dict(maskpadtype='randshift', cropmode='all'), dict(lengthmax=100, maskpadtype='padright'), dict(maskpadtype='randshift')]
        for i, kw in enumerate(cases):
            np.random.seed(40 + i)
            ids = [fids[j] for j in np.random.permutation(len(fids))]
            st = np.random.get_state()
            X, Y, W = data.load_inoutset(indir, outdir, wdir, ids, **kw)
            after = np.random.randint(0, 1 << 30)
            n = len(ids) // world
            for rank in range(world):
                np.random.set_state(st)
                Xs, Ys, Ws = data.load_inoutset_shard(indir, outdir, wdir, ids, rank, world, **kw)
                assert np.random.randint(0, 1 << 30) == after, (kw, rank)
                for got, want in ((Xs, X), (Ys, Y), (Ws, W)):
                    want = want[rank * n:(rank + 1) * n]
                    assert got.dtype == want.dtype and got.shape == want.shape and np.array_equal(got, want), (kw, rank)
        with pytest.raises(ValueError):
            data.load_inoutset_shard(indir, outdir, wdir, fids[:5], 0, 2)
        # a selector without a frame size: whole batch loaded, shard kept
        np.random.seed(3)
        X, Y, W = data.load_inoutset(indir, outdir, wdir, fids)
        np.random.seed(3)
        Xs, _, _ = data.load_inoutset_shard(indir.replace(':(-1,%d)' % S.CTX, ':(-1,dimfile)'), outdir, wdir, fids, 1, 2)
        assert np.array_equal(Xs, X[3:])
    finally:
        np.random.set_state(state)


def test_epoch_loop_feeds_the_batches_of_the_sequential_loop(tmp_path, monkeypatch):
    """train_oneparamset with its loader thread hands train_on_batch the same batches, in the same order, as a plain loop that
    shuffles and loads batch by batch from the same seed."""
    import torch
    from test_api_cpu import make_files, tiny_cfg
    from cpu_ops_double import CpuOpsDouble
    from percivaltts_b200 import backend_b200, modeltts_common, networks_critic, optimizertts_wgan, vocoders
    lens = [60, 64, 57, 53, 49, 56, 61, 50, 55, 58, 62, 51]
    fids, indir, outdir, wdir = make_files(tmp_path, lens)
    cfg = tiny_cfg()
    cfg.train_batch_size = 2
    cfg.train_min_nbepochs = 1
    cfg.train_max_nbepochs = 2
    cfg.train_cancel_nodecepochs = 5
    cfg.train_batch_lengthmax = None
    cfg.train_nbepochs_scalewdata = False
    backend_b200.set_ops(CpuOpsDouble(torch.float32))
    state = np.random.get_state()
    try:
        voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
        model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
        opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=networks_critic.Critic(voc, 425, cfg))
        opti.use_graphs = False
        seen = []

        def fake_train_on_batch(batchid, X, Y, next_batch=None):
            seen.append((batchid, X.copy(), Y.copy(), None if next_batch is None else next_batch[0].copy()))
            return 1.0
        monkeypatch.setattr(opti, 'train_on_batch', fake_train_on_batch)
        monkeypatch.setattr(opti, 'update_validation_cost', lambda costs, X_vals, Y_vals: (costs['model_rmse_validation'].append(1.0), 1.0)[1])
        tra, val = fids[:10], fids[10:]
        np.random.seed(99)
        opti.train(indir, outdir, wdir, tra, val, str(tmp_path / 'model.h5'))
        # the sequential loop of the reference, same seed
        np.random.seed(99)
        want = []
        nbb = len(tra) // cfg.train_batch_size
        for epoch in range(2):
            idx = np.arange(nbb * cfg.train_batch_size)
            np.random.shuffle(idx)
            for bid, part in enumerate(np.split(idx, nbb)):
                X, Y, W = data.load_inoutset(indir, outdir, wdir, [tra[i] for i in part], length=opti.cfg.train_batch_length,
                                             lengthmax=opti.cfg.train_batch_lengthmax, maskpadtype=opti.cfg.train_batch_padtype,
                                             cropmode=opti.cfg.train_batch_cropmode)
                want.append((bid, X, Y))
        assert len(seen) == len(want) == 2 * nbb
        for k, ((bid, X, Y, nxt), (wbid, wX, wY)) in enumerate(zip(seen, want)):
            assert bid == wbid and np.array_equal(X, wX) and np.array_equal(Y, wY), k
            if nxt is not None:
                assert np.array_equal(nxt, want[k + 1][1]), k
        assert sum(1 for s in seen if s[3] is None) <= 2          # only the last batch of an epoch has no successor to prefetch
    finally:
        np.random.set_state(state)
        backend_b200.set_ops(None)


def test_shard_loader_refuses_nan_in_a_window(tmp_path):
    import data_scenarios as S
    fids = S.make_files(str(tmp_path))
    indir = os.path.join(str(tmp_path), 'lab', '*.lab') + ':(-1,%d)' % S.CTX
    outdir = os.path.join(str(tmp_path), 'cmp', '*.cmp') + ':(-1,%d)' % S.OUT
    wdir = os.path.join(str(tmp_path), 'w', '*.w') + ':(-1,1)'
    f = os.path.join(str(tmp_path), 'cmp', fids[1] + '.cmp')
    Y = np.fromfile(f, dtype=np.float32)
    Y[S.OUT * 30 + 2] = np.nan
    Y.tofile(f)
    with pytest.raises(ValueError):
        data.load_inoutset_shard(indir, outdir, wdir, fids[:2], 0, 1)       # padright: every kept frame is in the window
