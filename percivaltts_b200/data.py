"""
Host-side batch assembly for the WGAN-GP step: raw float32 feature files -> cropped, windowed [B,T,feat] batches.
Same functions and semantics as reference percivaltts/data.py (path:shape suffix parser :34-76, load :96-133,
croplen :143-171, croplen_weight :173-231, batching :234-284, load_inoutset :297-322, cost helpers :327-406);
stays numpy on the host -- the device work starts at OptimizerTTSWGAN.train_on_batch.
"""
import os
import re
import time

import numpy as np

from .percivaltts import is_int, numpy_force_random_seed, print_tty

numpy_force_random_seed()      # reference data.py:32: repeatable shuffles / crops, and the SAME draws on every replica

_SHAPE_SUFFIX = re.compile(r'^(.*):\((.*)\)$')


def getpath(path):
    """Drop the optional ':(d0,d1,...)' shape suffix."""
    m = _SHAPE_SUFFIX.match(path)
    return m.group(1) if m else path


def getpathandshape(path, shape=None):
    """'/dir/*.spec:(-1,60)' -> ('/dir/*.spec', (-1, 60)).  A non-integer entry names a float32 file in the same
    directory whose length gives that dimension.  An explicit `shape` argument wins over the suffix."""
    m = _SHAPE_SUFFIX.match(path)
    if not m:
        return path, shape
    base, spec = m.group(1), m.group(2)
    dims = []
    for tok in spec.split(','):
        tok = tok.strip()
        if is_int(tok):
            dims.append(int(tok))
        else:
            ref = np.fromfile(os.path.join(os.path.dirname(base), tok), dtype=np.float32)
            dims.append(ref.shape[0])
    if shape is not None:
        print('WARNING: shape given both as argument ({}) and as path suffix ({}); using the argument'.format(shape, tuple(dims)))
        return base, shape
    return base, tuple(dims)


def getlastdim(path):
    _, shape = getpathandshape(path)
    return 1 if shape is None else shape[-1]


def _read(fpath, shape):
    if not os.path.isfile(fpath):
        raise ValueError('{} does not exists'.format(fpath))
    X = np.fromfile(fpath, dtype='float32')
    if shape is not None:
        X = X.reshape(shape)
    if not np.isfinite(X).all():
        raise ValueError('ERROR: There are nan or inf in {}'.format(fpath))
    return X


def loadfile(fpath, fbase=None, shape=None):
    if fbase is not None:
        fpath = fpath.replace('*', fbase)
    fpath, shape = getpathandshape(fpath, shape)
    return _read(fpath, shape)


def load(dirpath, fbases, shape=None, frameshift=0.005, verbose=0, label=''):
    """One matrix per file id, as a python list (lengths differ)."""
    dirpath, shape = getpathandshape(dirpath, shape)
    Xs, frames, mbytes = [], 0, 0.0
    for n, fbase in enumerate(fbases):
        if verbose > 0:
            print_tty('\r    {}Loading file {}/{} {}: ({:.2f}% done)        '.format(label, 1 + n, len(fbases), fbase,
                                                                               100.0 * n / len(fbases)))
        X = _read(dirpath.replace('*', fbase), shape)
        Xs.append(X)
        frames += X.shape[0]
        mbytes += X.size * 4.0 / (1024 ** 2)
    if verbose > 0:
        print_tty('\r' + ' ' * 64 + '\r')
        print('    {}{} sentences, frames={} ({}), {} MB'.format(label, len(fbases), frames,
                                                               time.strftime('%H:%M:%S', time.gmtime(frames * frameshift)), mbytes))
    return Xs


def gettotallen(Xs, axis=0):
    return int(sum(x.shape[axis] for x in Xs))


def _same_count(lists):
    if len(set(len(x) for x in lists)) > 1:
        raise ValueError('the size of the data sets are not identical ({})'.format([len(x) for x in lists]))


def croplen(xs, axis=0):
    """Trim sample k of every list in xs to the shortest of its versions along `axis` (lists modified in place)."""
    if axis > 2:
        raise ValueError('Do not manage axis values bigger than 2')
    _same_count(xs)
    for k in range(len(xs[0])):
        n = min(x[k].shape[axis] for x in xs)
        sl = (slice(None),) * axis + (slice(0, n),)
        for x in xs:
            x[k] = x[k][sl]
    return xs


def _speech_mask(w, thresh):
    return (w[:, 0] if w.ndim > 1 else w) > thresh


def _crop_selection(wk, thresh, cropmode, cropsize):
    """What croplen_weight keeps of one utterance with time weights wk: a slice or an index array."""
    keep = _speech_mask(wk, thresh)
    idx = np.where(keep)[0]
    if cropmode == 'begend':
        return slice(int(idx.min()), int(idx.max()))
    if cropmode == 'begendbigger':
        gaps = np.diff(idx)
        for g in np.where(gaps > 1)[0]:
            if gaps[g] < int(cropsize):
                keep[idx[g]:idx[g + 1]] = True
        return np.where(keep)[0]
    if cropmode == 'all':
        return idx
    raise ValueError('unknown cropmode ' + str(cropmode))


def croplen_weight(xs, w, thresh=0.5, cropmode='begend', cropsize=int(0.750 / 0.005)):
    """Remove silences according to the time weights w (threshold `thresh`):
    'begend' keeps [first speech frame, last speech frame); 'begendbigger' also drops inner pauses of at least
    `cropsize` frames; 'all' keeps speech frames only."""
    _same_count([w] + list(xs))
    for k in range(len(w)):
        sel = _crop_selection(w[k], thresh, cropmode, cropsize)
        for x in xs:
            x[k] = x[k][sel]
        w[k] = w[k][sel]
    return xs, w


def batching(xs, length=None, lengthmax=None, padtype='randshift', outmask=False):
    """Stack variable-length samples into [B, length, feat] float32 arrays.  'padright' zero-pads to the longest
    sample; 'randshift' takes a random window of the shortest length (one np.random.randint per sample, in sample
    order -- the draw order is part of the reference's repeatability contract, data.py:276-277)."""
    _same_count(xs)
    lens = [x.shape[0] for x in xs[0]]
    if length is None:
        length = max(lens) if padtype == 'padright' else min(lens)
    if lengthmax is not None and length > lengthmax:
        length = lengthmax
    length = int(length)
    nb = len(xs[0])
    outs = []
    for x in xs:
        feat = 1 if x[0].ndim == 1 else x[0].shape[1]
        outs.append(np.zeros((nb, length, feat), dtype='float32'))
    MB = np.zeros((nb, length), dtype='float32') if outmask else None
    shift = 0
    for b in range(nb):
        n = min(lens[b], length)
        if padtype == 'randshift':
            shift = np.random.randint(0, (lens[b] - length) + 1)
        for x, out in zip(xs, outs):
            out[b, :n, :] = x[b][shift:shift + n].reshape(n, -1)
        if outmask:
            MB[b, :n] = 1
    return outs, MB


def addstop(X, value=1.0):
    """A copy of the list with one more column (zeros) on every matrix and one more row [0, ..., 0, value] behind it: the stop
    symbol of sequences whose input and output lengths differ (data.py:286-295).  float64, as numpy promotes in the reference."""
    out = []
    for x in X:
        y = np.zeros((x.shape[0] + 1, x.shape[1] + 1), dtype=np.float64)
        y[:-1, :-1] = x
        y[-1, -1] = value
        out.append(y)
    return out


def load_inoutset(indir, outdir, outwdir, fid_lst, inouttimesync=True, length=None, lengthmax=None, maskpadtype='padright',
                  cropmode='begend', verbose=0):
    """Load inputs, outputs and time weights of the given ids, crop silences, and batch them."""
    X = load(indir, fid_lst, verbose=verbose, label='Context labels: ')
    Y = load(outdir, fid_lst, verbose=verbose, label='Output features: ')
    Wt = load(outwdir, fid_lst, verbose=verbose, label='Time weights: ')
    if not inouttimesync:
        # inputs and outputs of different lengths (data.py:309-313,319-321): a stop frame behind each, batched separately (two
        # rounds of window draws), the time weights stay a list of cropped matrices
        X = addstop(X)
        Y, Wt = croplen([Y, Wt])
        [Y], Wt = croplen_weight([Y], Wt, cropmode=cropmode)
        Y = addstop(Y)
        [X], _ = batching([X], length=length, lengthmax=lengthmax, padtype=maskpadtype)
        [Y], _ = batching([Y], length=length, lengthmax=lengthmax, padtype=maskpadtype)
        return X, Y, Wt
    X, Y, Wt = croplen([X, Y, Wt])
    [X, Y], Wt = croplen_weight([X, Y], Wt, cropmode=cropmode)
    [X, Y, Wt], _ = batching([X, Y, Wt], length=length, lengthmax=lengthmax, padtype=maskpadtype)
    return X, Y, Wt


def _rows_on_disk(path, fbase):
    """Frames of one raw float32 file whose selector is ':(-1,D)', from its size alone; None when the selector has another form."""
    fpath, shape = getpathandshape(path)
    if shape is None or len(shape) != 2 or shape[0] != -1 or shape[1] < 1:
        return None
    fpath = fpath.replace('*', fbase)
    if not os.path.isfile(fpath):
        raise ValueError('{} does not exists'.format(fpath))
    return os.path.getsize(fpath) // (4 * shape[1])


def load_inoutset_shard(indir, outdir, outwdir, fid_lst, rank, world, length=None, lengthmax=None, maskpadtype='padright',
                        cropmode='begend', verbose=0):
    """Utterances [rank*B/world, (rank+1)*B/world) of load_inoutset(...)'s batch -- the same arrays, the same draws from numpy's
    global RNG -- reading the input and output files of THOSE utterances only.  What has to be known of the whole batch (its
    common length and one window draw per utterance, in utterance order) follows from the file sizes and the time weights, which
    are one float per frame and are read for every utterance.  No reference counterpart (the reference is one process)."""
    nb = len(fid_lst)
    if nb % world:
        raise ValueError('batch of {} utterances does not split over {} ranks'.format(nb, world))
    nx = [_rows_on_disk(indir, f) for f in fid_lst]
    ny = [_rows_on_disk(outdir, f) for f in fid_lst]
    if any(n is None for n in nx + ny):           # a selector the sizes cannot be read from: load everything, keep the shard
        X, Y, Wt = load_inoutset(indir, outdir, outwdir, fid_lst, length=length, lengthmax=lengthmax, maskpadtype=maskpadtype,
                                 cropmode=cropmode, verbose=verbose)
        n = nb // world
        return X[rank * n:(rank + 1) * n], Y[rank * n:(rank + 1) * n], Wt[rank * n:(rank + 1) * n]
    Wt = load(outwdir, fid_lst, verbose=verbose, label='Time weights: ')
    xpat, ypat = [(fp, sh[1]) for fp, sh in (getpathandshape(indir), getpathandshape(outdir))]
    sels, lens = [], []
    for k in range(nb):
        n = min(nx[k], ny[k], Wt[k].shape[0])                                     # croplen
        Wt[k] = Wt[k][:n]
        sel = _crop_selection(Wt[k], 0.5, cropmode, int(0.750 / 0.005))           # croplen_weight
        Wt[k] = Wt[k][sel]
        sels.append((n, sel))                                                     # sel indexes the first n frames
        lens.append(Wt[k].shape[0])
    if length is None:                                                            # batching
        length = max(lens) if maskpadtype == 'padright' else min(lens)
    if lengthmax is not None and length > lengthmax:
        length = lengthmax
    length = int(length)
    per = nb // world
    mine = range(rank * per, (rank + 1) * per)
    outs, reads = None, []
    for b in range(nb):
        shift = np.random.randint(0, (lens[b] - length) + 1) if maskpadtype == 'randshift' else 0
        if b not in mine:
            continue
        m = min(lens[b], length)
        sel = sels[b][1]
        if outs is None:      # every frame of every row is written below (window, then zeros behind a short utterance): no zero fill
            outs = [np.empty((per, length, d), dtype='float32') for d in (xpat[1], ypat[1], 1 if Wt[b].ndim == 1 else Wt[b].shape[1])]
        r = b - mine[0]
        # only the window's frames leave the page cache, read straight into the batch: rows [shift, shift + m) of the cropped
        # utterance (the whole-file read, crop copy and window copy of load_inoutset are three passes more)
        reads.append((outs[0][r, :m], xpat, fid_lst[b], sel, shift, m))
        reads.append((outs[1][r, :m], ypat, fid_lst[b], sel, shift, m))
        outs[2][r, :m] = Wt[b][shift:shift + m].reshape(m, -1)
        for out in outs:
            out[r, m:] = 0
    # every window draw is made (in utterance order, above); the reads themselves are independent and release the GIL
    if len(reads) > 2:
        list(_read_pool().map(lambda a: _read_window_into(*a), reads))
    else:
        for a in reads:
            _read_window_into(*a)
    for out, path in zip(outs, (indir, outdir, outwdir)):
        # a NaN or an Inf anywhere makes the plain sum non-finite; so can an overflow of finite values, hence the exact second look
        if not np.isfinite(out.sum()) and not np.isfinite(out).all():
            raise ValueError('ERROR: There are nan or inf in a file of {}'.format(getpath(path)))
    return outs[0], outs[1], outs[2]


_READ_POOL = None


def _read_pool():
    """Four threads shared by every batch assembly of the process (file reads into disjoint rows of a batch)."""
    global _READ_POOL
    if _READ_POOL is None:
        import concurrent.futures
        _READ_POOL = concurrent.futures.ThreadPoolExecutor(max_workers=4, thread_name_prefix='pctts-read')
    return _READ_POOL


def _read_window_into(dst, pat, fbase, sel, shift, m):
    """Frames [shift, shift + m) of what the crop selection `sel` keeps of one ':(-1,D)' file -> dst [m, D] (C-contiguous view of
    the batch).  Consecutive frames (always for 'begend'; for the other modes unless a dropped pause falls inside the window) are one
    seek + one read into dst; otherwise the span is read once and the kept frames gathered."""
    fpath, D = pat
    if m <= 0:
        return
    if isinstance(sel, slice):
        first, idx = sel.start + shift, None
    else:
        idx = sel[shift:shift + m]
        first = int(idx[0])
        if int(idx[-1]) - first + 1 == m:
            idx = None
    with open(fpath.replace('*', fbase), 'rb', buffering=0) as f:
        f.seek(first * D * 4)
        if idx is None:
            if f.readinto(memoryview(dst).cast('B')) != m * D * 4:
                raise ValueError('{} is shorter than its time weights say'.format(fpath.replace('*', fbase)))
        else:
            span = int(idx[-1]) - first + 1
            dst[...] = np.fromfile(f, dtype=np.float32, count=span * D).reshape(span, D)[idx - first]


# ---- evaluation helpers ---------------------------------------------------------------------------------------------

def cost_0pred_rmse(Y_val):
    """RMSE of the all-zero predictor (the data's own RMS)."""
    if isinstance(Y_val, list):
        tot = sum(float(np.sum(np.square(y, dtype=np.float64))) for y in Y_val)
        return float(np.sqrt(tot / sum(y.size for y in Y_val)))
    return float(np.sqrt(np.mean(np.square(Y_val))))


def _one_by_one(Xs):
    for k in range(len(Xs[0])):
        yield k, [np.asarray(inp[k])[np.newaxis] for inp in Xs]


def cost_model_mfn(fn, Xs):
    """Mean over utterances of fn(sample 0 of every list, ...), one utterance at a time."""
    if not isinstance(Xs[0], list):
        return 0.0
    return sum(fn(*ins) for _, ins in _one_by_one(Xs)) / len(Xs[0])


def cost_model_prediction_rmse(mod, Xs, Y_val, inouttimesync=True):
    if not isinstance(Xs[0], list):
        return 0.0
    sq, n = 0.0, 0
    for k, ins in _one_by_one(Xs):
        yp = mod.predict(*ins)[0]
        sq += float(np.sum((Y_val[k] - yp) ** 2))
        n += yp.size
    return float(np.sqrt(sq / n))


def prediction_rms(mod, Xs):
    if not isinstance(Xs[0], list):
        return 0.0
    sq, n = 0.0, 0
    for _, ins in _one_by_one(Xs):
        yp = mod.predict(*ins)[0]
        sq += float(np.sum(yp ** 2))
        n += yp.size
    return float(np.sqrt(sq / n))


def prediction_mstd(mod, Xs):
    if not isinstance(Xs[0], list):
        return 0.0
    return float(np.mean([np.std(mod.predict(*ins)[0]) for _, ins in _one_by_one(Xs)]))
