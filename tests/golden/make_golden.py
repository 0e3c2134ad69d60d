"""
Generates the committed fixtures under tests/golden/.  Run in the build container (where /root/reference exists):
    python tests/golden/make_golden.py

1. slt_arctic_label_durations.json -- per-utterance HTS state durations (100 ns units) read from the reference's
   bundled fixture tests/slt_arctic_merlin_test.tar.gz (label_state_align/*.lab).  The reference's own test pins
   the resulting frame counts: Xs[0].shape==(667,425) and gettotallen(Xs)==5694 (tests/test_smoke.py:80-85).
2. pml_golden.npz -- oracle/pml_oracle.py outputs (float64) on seeded inputs: regression pins of the operator
   DEFINITION (the reference's sigproc arithmetic is unavailable: parity unpinned, see oracle/pml_oracle.py).
3. step_golden.json -- oracle/percival_oracle.py losses and gradient checksums (float64) of the tiny configuration
   (tests/test_smoke_tensorflowkeras.py:183-190): regression pins of the oracle itself.
"""
import io
import json
import os
import sys
import tarfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def label_durations():
    tar = '/root/reference/tests/slt_arctic_merlin_test.tar.gz'
    out = {}
    with tarfile.open(tar) as tf:
        ids = tf.extractfile('slt_arctic_merlin_test/file_id_list.scp').read().decode().split()
        for fid in ids:
            txt = tf.extractfile('slt_arctic_merlin_test/label_state_align/%s.lab' % fid).read().decode()
            durs = []
            for ln in txt.splitlines():
                parts = ln.split()
                if len(parts) >= 3:
                    durs.append(int(parts[1]) - int(parts[0]))
            out[fid] = durs
    with open(os.path.join(HERE, 'slt_arctic_label_durations.json'), 'w') as f:
        json.dump({'ids': ids, 'durations_100ns': out}, f)
    print('label durations:', [sum(int(d / 50000) for d in out[i]) for i in ids])


def pml_golden():
    from oracle import pml_oracle as P
    rng = np.random.RandomState(123)
    T, fs, dftlen = 6, 16000, 4096
    nb = dftlen // 2 + 1
    SPEC = np.exp(rng.normal(-6, 2, size=(T, nb))).astype(np.float32)
    NM = np.clip(rng.uniform(-0.5, 1.5, size=(T, nb)), 0, 1).astype(np.float32)
    d = dict(SPEC=SPEC, NM=NM)
    for K in (129, 65):
        d['cspec%d' % K] = P.compress_spectrum(SPEC, fs, K)
        d['dspec%d' % K] = P.decompress_spectrum(d['cspec%d' % K].astype(np.float32), fs, dftlen)
    for K in (33, 17):
        d['cnm%d' % K] = P.linbnd2fwbnd(NM, fs, dftlen, K)
        d['dnm%d' % K] = P.fwbnd2linbnd(d['cnm%d' % K].astype(np.float32), fs, dftlen)
    d['fwidx'] = np.array([P.freq2fwspecidx(4000, 16000, 129), P.freq2fwspecidx(4000, 16000, 65),
                           P.freq2fwspecidx(0, 16000, 129), P.freq2fwspecidx(8000, 16000, 129)])
    np.savez_compressed(os.path.join(HERE, 'pml_golden.npz'), **d)
    print('pml golden:', sorted(d))


def step_golden():
    import torch
    from oracle import percival_oracle as O
    a = O.tiny_arch()
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(a), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(a), seed=2))
    X, Y, alpha = O.synthetic_batch(a, 2, 29, seed=5)
    dt = torch.float64
    wgt, wct = O.to_torch(wg, dt, True), O.to_torch(wc, dt, True)
    Xt, Yt = torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt)
    rc = O.critic_step(a, wgt, wct, Xt, Yt, torch.tensor(alpha, dtype=dt).reshape(-1, 1, 1))
    rg = O.generator_step(a, wgt, wct, Xt, Yt)
    out = dict(critic_loss=float(rc['loss']), l_real=float(rc['l_real']), l_fake=float(rc['l_fake']), gp=float(rc['gp']),
               gen_loss=float(rg['loss']), l_w=float(rg['l_w']), l_ls=float(rg['l_ls']),
               critic_grad_l1=dict((k, float(v.abs().sum())) for k, v in rc['grads'].items()),
               gen_grad_l1=dict((k, float(v.abs().sum())) for k, v in rg['grads'].items()),
               pred_l1=float(rg['pred'].abs().sum()))
    with open(os.path.join(HERE, 'step_golden.json'), 'w') as f:
        json.dump(out, f, indent=1)
    print('step golden: critic %.6f gen %.6f' % (out['critic_loss'], out['gen_loss']))


if __name__ == '__main__':
    label_durations()
    pml_golden()
    step_golden()
