"""
Frequency-warped band compression of PML spectral envelopes / noise masks on the B200:
`linbnd2fwbnd`, `fwbnd2linbnd`, `freq2fwspecidx` -- the three functions of the reference's `sigproc` dependency
(gitlab.com/gillesdegottex/sigproc @ 289d9dff, absent from the reference tree) that sit on the hot path.

Call sites in the reference: percivaltts/vocoders.py:139,150,202; pulsemodel/analysis.py:336,366;
pulsemodel/synthesis.py:316,351; percivaltts/optimizertts_wgan.py:192,200.

Operator definition (sigproc's source is unavailable, so it is fixed here and in oracle/pml_oracle.py from what the
call sites constrain -- mel bands including the DC and Nyquist ends, convex interpolation, a smooth variant):
  mel(f) = 1127 ln(1 + f/700);  band k at f_k = mel^-1(k/(K-1) mel(fs/2))
  linbnd2fwbnd : linear interpolation of each row (bins at i fs/dftlen) at the f_k
  fwbnd2linbnd : linear (or Catmull-Rom when smooth=True) interpolation of each row of bands at
                 u_i = (K-1) mel(i fs/dftlen)/mel(fs/2)
The interpolation tables (index + weight per output column) are built here in float64 and are *inputs* of the CUDA
kernels (pctts_pml_linbnd2fwbnd / pctts_pml_fwbnd2linbnd), so other warpings can be substituted without new kernels.
"""
import numpy as np
import torch

from . import backend_b200


def mel(f):
    return 1127.0 * np.log1p(np.asarray(f, dtype=np.float64) / 700.0)


def imel(m):
    return 700.0 * np.expm1(np.asarray(m, dtype=np.float64) / 1127.0)


def freq2fwspecidx(freq, fs, nbbnds):
    """Fractional band index of `freq` on the warped axis (used by the LSE/WGAN loss weighting)."""
    return (nbbnds - 1) * mel(freq) / mel(0.5 * fs)


def compress_table(fs, dftlen, nbbnds):
    """-> (idx int32 [K], w float32 [K]): band k = (1-w) x[idx] + w x[idx+1]."""
    nbins = dftlen // 2 + 1
    k = np.arange(nbbnds, dtype=np.float64)
    p = imel(k / (nbbnds - 1) * mel(0.5 * fs)) * dftlen / float(fs)
    p[0], p[-1] = 0.0, nbins - 1
    idx = np.minimum(np.floor(p), nbins - 1)
    return idx.astype(np.int32), (p - idx).astype(np.float32)


def decompress_table(fs, dftlen, nbbnds):
    """-> (idx int32 [nbins], w float32 [nbins]): bin i interpolates the bands at idx[i] + w[i]."""
    i = np.arange(dftlen // 2 + 1, dtype=np.float64)
    u = (nbbnds - 1) * mel(i * fs / float(dftlen)) / mel(0.5 * fs)
    u[0], u[-1] = 0.0, nbbnds - 1
    idx = np.minimum(np.floor(u), nbbnds - 2)
    return idx.astype(np.int32), (u - idx).astype(np.float32)


_tables = {}


def _device_table(kind, fs, dftlen, nbbnds, ops):
    key = (kind, float(fs), int(dftlen), int(nbbnds), str(ops.device))
    t = _tables.get(key)
    if t is None:
        idx, w = (compress_table if kind == 'c' else decompress_table)(fs, dftlen, nbbnds)
        t = _tables[key] = (torch.from_numpy(idx).to(ops.device), torch.from_numpy(w).to(ops.device))
    return t


def linbnd2fwbnd_device(x, fs, dftlen, nbbnds, log_abs=False, out=None, ops=None):
    """x: device float32 [T, dftlen/2+1] -> device float32 [T, nbbnds] (log|x| interpolated when log_abs)."""
    ops = ops or backend_b200.get_ops()
    assert x.shape[1] == dftlen // 2 + 1
    idx, w = _device_table('c', fs, dftlen, nbbnds, ops)
    if out is None:
        out = ops.empty(x.shape[0], nbbnds)
    ops.pml_linbnd2fwbnd(x, idx, w, out, log_abs)
    return out


def fwbnd2linbnd_device(y, fs, dftlen, smooth=False, do_exp=False, out=None, ops=None):
    """y: device float32 [T, nbbnds] -> device float32 [T, dftlen/2+1]."""
    ops = ops or backend_b200.get_ops()
    idx, w = _device_table('d', fs, dftlen, y.shape[1], ops)
    if out is None:
        out = ops.empty(y.shape[0], dftlen // 2 + 1)
    ops.pml_fwbnd2linbnd(y, idx, w, out, smooth, do_exp)
    return out


def _to_device(X, ops):
    return torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32)).to(ops.device)


def linbnd2fwbnd(X, fs, dftlen, nbbnds):
    """sp.linbnd2fwbnd: numpy [T, dftlen/2+1] -> numpy float32 [T, nbbnds]."""
    ops = backend_b200.get_ops()
    return linbnd2fwbnd_device(_to_device(X, ops), fs, dftlen, nbbnds, ops=ops).cpu().numpy()


def fwbnd2linbnd(Y, fs, dftlen, smooth=False):
    """sp.fwbnd2linbnd: numpy [T, nbbnds] -> numpy float32 [T, dftlen/2+1]."""
    ops = backend_b200.get_ops()
    return fwbnd2linbnd_device(_to_device(Y, ops), fs, dftlen, smooth=smooth, ops=ops).cpu().numpy()
