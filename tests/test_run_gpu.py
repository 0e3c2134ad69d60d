"""
BASELINE config #1 on the GPU: the reference's tiny WLSWGAN smoke run (tests/test_smoke_tensorflowkeras.py:183-203) through
the public API on the real CUDA backend -- 8 training + 2 validation utterances whose frame counts are those of the bundled
slt_arctic_merlin_test labels (667, 747, ... frames; features synthetic, the reference's analysis chain is out of scope),
batch size 2, CUDA-graph trainer, checkpoint + resume, generation of .cmp files and the device-side PML decompression of the
generated features (vocoders.py:194-202).  Purpose, as in the reference: the loop runs, costs are finite, files are written.
"""
import os

import numpy as np
import pytest
import torch

from percivaltts_b200 import backend_b200, modeltts_common, networks_critic, optimizertts_wgan, vocoders
from test_api_cpu import make_files, tiny_cfg

pytestmark = pytest.mark.gpu

LABEL_FRAMES = [667, 747, 637, 497, 293, 589, 597, 453, 615, 599]      # tests/test_smoke.py:80-85, re-derived from the .lab files


def test_tiny_wlswgan_run_on_gpu(cuda_ops, tmp_path):
    backend_b200.set_ops(cuda_ops)
    try:
        lens = [n - 1 for n in LABEL_FRAMES]
        fids, indir, outdir, wdir = make_files(tmp_path, lens)
        cfg = tiny_cfg()
        cfg.train_min_nbepochs = 1
        cfg.train_max_nbepochs = 3
        cfg.train_cancel_nodecepochs = 3
        cfg.train_batch_lengthmax = 200          # keeps the number of captured (B, T) shapes small
        cfg.train_nbepochs_scalewdata = False
        voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
        model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
        critic = networks_critic.Critic(voc, 425, cfg)
        opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=critic)
        fmodel = str(tmp_path / 'model.h5')
        n0 = cuda_ops.lib.pctts_launch_count()
        res = opti.train(indir, outdir, wdir, fids[:8], fids[8:], fmodel)
        assert cuda_ops.lib.pctts_launch_count() > n0            # the kernels of libpctts.so did the work
        assert res['epoch_stopped'] == 3 and np.isfinite(res['best_val'])
        for suffix in ('.arch.json', '.weights.npz', '.cfgextras.pkl'):
            assert os.path.exists(fmodel + suffix)
        # save -> load round trip (tests/test_smoke_tensorflowkeras.py:77-79): the reloaded model predicts what the trained one does
        model2 = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
        cfg_loaded, extras = model2.load(fmodel)
        assert 'cost_val' in extras
        x = np.random.RandomState(1).uniform(-1, 1, size=(1, 321, 425)).astype(np.float32)
        y1, y2 = model.predict(x), model2.predict(x)
        assert y1.shape == (1, 321, 83) and np.isfinite(y1).all() and np.array_equal(y1, y2)
        # resume for one more epoch
        opti.cfg.train_max_nbepochs = 4
        res = opti.train(indir, outdir, wdir, fids[:8], fids[8:], fmodel, cont=True)
        assert res['epoch_stopped'] == 4
        # generation tail: .cmp files, then f0 / spectral envelope / noise mask on the linear frequency axis
        model.generate_cmp(indir, str(tmp_path / 'out' / '*.cmp'), fids[8:])
        gen = np.fromfile(str(tmp_path / 'out' / (fids[8] + '.cmp')), dtype=np.float32).reshape(-1, 83)
        assert gen.shape[0] == LABEL_FRAMES[8] and np.isfinite(gen).all()
        f0, SPEC, NM = voc.decompress_features(gen)
        assert f0.shape == (gen.shape[0],) and SPEC.shape == (gen.shape[0], 2049) and NM.shape == (gen.shape[0], 2049)
        assert np.isfinite(SPEC).all() and (SPEC > 0).all()
        assert NM.min() >= 0.0 and NM.max() <= 1.0               # sigmoid outputs, convex interpolation
    finally:
        backend_b200.set_ops(None)
