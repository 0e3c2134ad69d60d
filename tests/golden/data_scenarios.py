"""
One scenario driver for the host-side batch assembly, run twice: by tests/golden/make_data_golden.py on the REFERENCE's own
percivaltts/data.py + percivaltts.py (imported from /root/reference in the build container; outputs committed as
tests/golden/data_golden.npz) and by tests/test_data_golden_cpu.py on percivaltts_b200.data / percivaltts_b200.percivaltts.
Everything the scenario reads is generated here from fixed seeds, so the GPU box and CI need neither the reference nor files.

run(data, pt, workdir) -> {name: ndarray}; arrays are compared bit for bit, names starting with 'cost.' within 1e-6 relative
(the reference accumulates float32 partial sums).
"""
import os

import numpy as np

LENS = [66, 74, 63, 49, 29, 58]
CTX, OUT = 7, 5


def make_files(workdir):
    rng = np.random.RandomState(20181)
    fids = ['utt%02d' % i for i in range(len(LENS))]
    for d in ('lab', 'cmp', 'w'):
        os.makedirs(os.path.join(workdir, d), exist_ok=True)
    for k, (fid, n) in enumerate(zip(fids, LENS)):
        rng.uniform(-1, 1, size=(n + 1 + (k % 3), CTX)).astype(np.float32).tofile(os.path.join(workdir, 'lab', fid + '.lab'))
        rng.normal(size=(n, OUT)).astype(np.float32).tofile(os.path.join(workdir, 'cmp', fid + '.cmp'))
        w = np.ones((n + (k % 2), 1), dtype=np.float32)
        w[:2 + k] = 0                                   # leading silence
        w[-(1 + k):] = 0                                # trailing silence
        w[20:20 + 3] = 0.25                             # a short pause (shorter than cropsize below)
        if n > 45:
            w[30:30 + 9] = 0                            # a long pause
        w.tofile(os.path.join(workdir, 'w', fid + '.w'))
    np.arange(CTX, dtype=np.float32).tofile(os.path.join(workdir, 'lab', 'dimfile'))     # a "shape by file length" selector
    with open(os.path.join(workdir, 'cfg_a.py'), 'w') as f:
        f.write('train_batch_size = 5\nid_valid_start = 23\nname = "a"\n')
    with open(os.path.join(workdir, 'cfg_b.py'), 'w') as f:
        f.write('id_valid_start = 48\nextra = [1, 2, 3]\n')
    with open(os.path.join(workdir, 'ids.scp'), 'w') as f:
        f.write('\n'.join(fids[:4]) + '\n')
    return fids


class LinearModel(object):
    """Stands in for a trained model: predict(X [1,T,CTX]) -> [1,T,OUT], a fixed linear map in float32."""

    def __init__(self):
        self.A = (np.arange(CTX * OUT, dtype=np.float32).reshape(CTX, OUT) - 17.0) / 23.0

    def predict(self, X):
        return np.matmul(X.astype(np.float32), self.A)


def _put(res, name, val):
    if isinstance(val, (list, tuple)) and len(val) and isinstance(val[0], np.ndarray):
        res[name + '.n'] = np.array(len(val))
        for i, v in enumerate(val):
            res['%s.%d' % (name, i)] = np.asarray(v)
    else:
        res[name] = np.asarray(val)


def _draw(res, name):
    """One more draw from numpy's global RNG: pins HOW MANY draws the call before it consumed."""
    res[name + '.nextdraw'] = np.array(np.random.randint(0, 1 << 30))


def run(data, pt, workdir):
    res = {}
    fids = make_files(workdir)
    lab = os.path.join(workdir, 'lab', '*.lab')
    indir, outdir, wdir = lab + ':(-1,%d)' % CTX, os.path.join(workdir, 'cmp', '*.cmp') + ':(-1,%d)' % OUT, \
        os.path.join(workdir, 'w', '*.w') + ':(-1,1)'

    # -- path:shape selectors (data.py:34-76) --
    _put(res, 'getpath', np.array([data.getpath(indir) == lab, data.getpath(lab) == lab]))
    p, s = data.getpathandshape(indir)
    _put(res, 'pathshape.int', np.array(list(s)))
    p, s = data.getpathandshape(lab + ':(-1,dimfile)')
    _put(res, 'pathshape.file', np.array(list(s)))
    p, s = data.getpathandshape(indir, shape=(-1, 1))
    _put(res, 'pathshape.argwins', np.array(list(s)))
    p, s = data.getpathandshape(lab)
    _put(res, 'pathshape.none', np.array([s is None, p == lab]))
    _put(res, 'lastdim', np.array([data.getlastdim(indir), data.getlastdim(lab), data.getlastdim(lab + ':(-1,dimfile)')]))

    # -- loading (data.py:78-141) --
    _put(res, 'loadfile', data.loadfile(outdir, fids[2]))
    _put(res, 'loadfile.flat', data.loadfile(os.path.join(workdir, 'cmp', fids[1] + '.cmp')))
    _put(res, 'loadfile.shapearg', data.loadfile(os.path.join(workdir, 'cmp', '*.cmp'), fids[1], shape=(-1, OUT)))
    X = data.load(indir, fids)
    Y = data.load(outdir, fids)
    W = data.load(wdir, fids)
    _put(res, 'load.X', X)
    _put(res, 'totallen', np.array([data.gettotallen(X), data.gettotallen(Y), data.gettotallen(W)]))

    # -- croplen (data.py:143-171) --
    X, Y, W = data.croplen([X, Y, W])
    _put(res, 'croplen.X', X)
    _put(res, 'croplen.W', W)

    # -- croplen_weight, the three crop modes (data.py:173-231) --
    for mode in ('begend', 'begendbigger', 'all'):
        Xc, Yc, Wc = [x.copy() for x in X], [y.copy() for y in Y], [w.copy() for w in W]
        [Xc, Yc], Wc = data.croplen_weight([Xc, Yc], Wc, cropmode=mode, cropsize=6)
        _put(res, 'cropw.%s.X' % mode, Xc)
        _put(res, 'cropw.%s.Y' % mode, Yc)
        _put(res, 'cropw.%s.W' % mode, Wc)
    Wflat = [w[:, 0].copy() for w in W]                                   # 1-D weights take the other branch
    Xc = [x.copy() for x in X]
    [Xc], Wflat = data.croplen_weight([Xc], Wflat, thresh=0.2, cropmode='begend')
    _put(res, 'cropw.flat.X', Xc)
    _put(res, 'cropw.flat.W', Wflat)

    # -- batching (data.py:234-284): window draws are one randint per sample, in sample order --
    cases = [('randshift', None, None), ('randshift', None, 20), ('randshift', 25, None), ('randshift', 40, 27),
             ('padright', None, None), ('padright', None, 50), ('padright', 31, None), ('other', None, None)]
    for i, (pad, length, lmax) in enumerate(cases):
        for outmask in (False, True):
            np.random.seed(100 + i)
            Xc, Yc, Wc = [x.copy() for x in X], [y.copy() for y in Y], [w[:, 0].copy() if pad == 'other' else w.copy() for w in W]
            if pad == 'other':
                [Xb, Yb], MB = data.batching([Xc, Yc], length=length, lengthmax=lmax, padtype=pad, outmask=outmask)
            else:
                [Xb, Yb, Wb], MB = data.batching([Xc, Yc, Wc], length=length, lengthmax=lmax, padtype=pad, outmask=outmask)
                _put(res, 'batch.%d.%d.W' % (i, outmask), Wb)
            _put(res, 'batch.%d.%d.X' % (i, outmask), Xb)
            _put(res, 'batch.%d.%d.Y' % (i, outmask), Yb)
            _put(res, 'batch.%d.%d.M' % (i, outmask), np.zeros(0) if MB is None else MB)
            _draw(res, 'batch.%d.%d' % (i, outmask))

    # -- load_inoutset (data.py:297-322), the call of the training loop (optimizertts.py:231-232) --
    sets = [dict(), dict(lengthmax=20, maskpadtype='randshift'), dict(length=16, maskpadtype='randshift', cropmode='begendbigger'),
            dict(maskpadtype='randshift', cropmode='all'), dict(lengthmax=100, maskpadtype='padright', cropmode='begend')]
    np.random.seed(123)
    for i, kw in enumerate(sets):
        for rep in range(2):                                               # consecutive batches continue the same RNG stream
            ids = [fids[j] for j in np.random.permutation(len(fids))[:4]]
            Xb, Yb, Wb = data.load_inoutset(indir, outdir, wdir, ids, **kw)
            _put(res, 'inout.%d.%d.X' % (i, rep), Xb)
            _put(res, 'inout.%d.%d.Y' % (i, rep), Yb)
            _put(res, 'inout.%d.%d.W' % (i, rep), Wb)
    _draw(res, 'inout')
    np.random.seed(7)
    Xb, Yb, Wb = data.load_inoutset(indir, outdir, wdir, fids[:3], inouttimesync=False, maskpadtype='randshift')
    _put(res, 'inout.nosync.X', Xb)
    _put(res, 'inout.nosync.Y', Yb)
    _put(res, 'inout.nosync.W', Wb if isinstance(Wb, np.ndarray) else list(Wb))
    _draw(res, 'inout.nosync')
    _put(res, 'addstop', data.addstop([x.copy() for x in X[:2]], value=0.5))

    # -- validation costs (data.py:327-406) --
    mod = LinearModel()
    _put(res, 'cost.0pred.list', data.cost_0pred_rmse(Y))
    _put(res, 'cost.0pred.array', data.cost_0pred_rmse(Yb))
    _put(res, 'cost.mfn', data.cost_model_mfn(lambda x, y: float(np.mean(x)) * 3.0 + float(np.mean(y * y)), [X, Y]))
    _put(res, 'cost.mfn.notlist', data.cost_model_mfn(lambda x: 1.0, [np.zeros((2, 3))]))
    _put(res, 'cost.rmse', data.cost_model_prediction_rmse(mod, [X], Y))
    _put(res, 'cost.mstd', data.prediction_mstd(mod, [X]))
    _put(res, 'cost.rms', data.prediction_rms(mod, [X]))

    # -- configuration object and small helpers (percivaltts.py:51-142) --
    cfg = pt.configuration()
    cfg.mergefiles(os.path.join(workdir, 'cfg_a.py'))
    a_nb = cfg.id_train_nb()
    cfg2 = pt.configuration()
    cfg2.mergefiles([os.path.join(workdir, 'cfg_a.py'), os.path.join(workdir, 'cfg_b.py')])
    other = pt.configuration()
    other.train_batch_size = 7
    other.fresh = 1.5
    cfg2.merge(other)
    _put(res, 'cfg', np.array([a_nb, cfg2.id_train_nb(), cfg2.id_valid_start, cfg2.train_batch_size, len(cfg2.extra),
                               int(cfg2.name == 'a'), int(cfg == cfg2), int(cfg != cfg2), int(cfg2.fresh * 2)]))
    same = pt.configuration()
    same.mergefiles(os.path.join(workdir, 'cfg_a.py'))
    _put(res, 'cfg.eq', np.array([int(cfg.train_batch_size == same.train_batch_size and cfg.id_valid_start == same.id_valid_start)]))
    _put(res, 'is_int', np.array([int(pt.is_int(v)) for v in ('12', '-3', 'a1', '1.5', '', ' 7 ', '1.0', '10', '100.', '0', '+4', '1..0', 5, -1)]))
    _put(res, 'readids', np.array([int(list(pt.readids(os.path.join(workdir, 'ids.scp'))) == fids[:4])]))
    _put(res, 'time2str', np.frombuffer('|'.join(pt.time2str(s) for s in (0, 59, 3600, 86399, 90061, 200000.7)).encode(), dtype=np.uint8))
    return res
