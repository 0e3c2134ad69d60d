"""
The hot path as the reference's own loop drives it: OptimizerTTSWGAN.train() on FILES (reference optimizertts.py:149-303 -- shuffle,
load_inoutset per batch from disk, crop silences, random windows, train_on_batch, validation and checkpoint every epoch), demo
architecture, B=10 utterances per batch cropped to T=400.  Synthetic float32 files are written to a scratch directory first.
Reported: frames/s over the batch loops of the epochs after the first (the first one warms up and captures the step graphs), with
the time the loop waited for data beside the time inside train_on_batch -- the reference prints the same two sums per epoch
(optimizertts.py:261).  Used by bench.py (key `e2e_files`) and runnable alone: python tools/bench_files.py
"""
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np


def write_corpus(workdir, n_utt, ctx, out, rng, lo=430, hi=620):
    fids = ['utt%03d' % i for i in range(n_utt)]
    for d in ('lab', 'cmp', 'w'):
        os.makedirs(os.path.join(workdir, d), exist_ok=True)
    for fid in fids:
        n = int(rng.randint(lo, hi))
        rng.uniform(-1, 1, size=(n + 1, ctx)).astype(np.float32).tofile(os.path.join(workdir, 'lab', fid + '.lab'))
        Y = rng.normal(size=(n, out)).astype(np.float32)
        Y[:, -33:] = rng.uniform(0, 1, size=(n, 33))
        Y.tofile(os.path.join(workdir, 'cmp', fid + '.cmp'))
        w = np.ones((n, 1), dtype=np.float32)
        w[:6] = 0
        w[-6:] = 0
        w.tofile(os.path.join(workdir, 'w', fid + '.w'))
    return (fids, os.path.join(workdir, 'lab', '*.lab') + ':(-1,%d)' % ctx, os.path.join(workdir, 'cmp', '*.cmp') + ':(-1,%d)' % out,
            os.path.join(workdir, 'w', '*.w') + ':(-1,1)')


def run(cfg, epochs=7, n_tra=80, n_val=2):
    from percivaltts_b200 import modeltts_common, networks_critic, optimizertts_wgan, vocoders
    workdir = tempfile.mkdtemp(prefix='pctts_bench_files_')
    stdout = sys.stdout
    try:
        rng = np.random.RandomState(5)
        voc = vocoders.VocoderPML(16000, 0.005, 129, 33)
        fids, indir, outdir, wdir = write_corpus(workdir, n_tra + n_val, 425, voc.featuressize(), rng)
        cfg.train_min_nbepochs = 1
        cfg.train_max_nbepochs = epochs
        cfg.train_cancel_nodecepochs = epochs + 1
        cfg.train_nbepochs_scalewdata = False
        sys.stdout = open(os.devnull, 'w')
        model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
        critic = networks_critic.Critic(voc, 425, cfg)
        opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=critic)
        t0 = time.perf_counter()
        opti.train(indir, outdir, wdir, fids[:n_tra], fids[n_tra:], os.path.join(workdir, 'model.h5'))
        total = time.perf_counter() - t0
        sys.stdout = stdout
        tim = opti.epoch_timings[1:]
        frames = sum(t['frames'] for t in tim)
        wait, train = sum(t['wait_for_data_s'] for t in tim), sum(t['train_s'] for t in tim)
        nb = sum(t['batches'] for t in tim)
        return {'value': frames / (wait + train), 'unit': 'frames/s', 'epochs_timed': len(tim), 'batches_timed': nb,
                'frames_per_batch': frames // max(nb, 1), 'ms_per_batch': 1e3 * (wait + train) / max(nb, 1),
                'ms_per_batch_waiting_for_data': 1e3 * wait / max(nb, 1), 'ms_per_batch_in_train_on_batch': 1e3 * train / max(nb, 1),
                'whole_train_call_s': total,
                'call': 'OptimizerTTSWGAN.train(indir, outdir, wdir, ids_train, ids_valid, model.h5) on %d + %d synthetic float32 files '
                        '(scratch directory, page cache warm): shuffle + load_inoutset + silence crop + random 400-frame windows per '
                        'batch on a loader thread two batches ahead, train_on_batch with the next batch prefetched; batch loops of '
                        'epochs 2..%d (validation and checkpoint writes of every epoch are in whole_train_call_s only)' % (n_tra, n_val, epochs)}
    finally:
        sys.stdout = stdout
        shutil.rmtree(workdir, ignore_errors=True)


if __name__ == '__main__':
    sys.path.insert(0, ROOT)
    import bench
    print(json.dumps(run(bench.demo_cfg())))
