"""
CPU ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by the product package).

PML band compression / decompression:  sp.linbnd2fwbnd, sp.fwbnd2linbnd, sp.freq2fwspecidx.

PARITY UNPINNED.  The arithmetic lives in the third-party `sigproc` package
(https://gitlab.com/gillesdegottex/sigproc.git @ 289d9dffe1f6ba84fa1fe962d6fb6e1062ec1cf6, see the
reference's .SUBMODULES.json:42-48) which is NOT present under /root/reference (empty directory
percivaltts/external/pulsemodel/sigproc/) and the reference's tests at that boundary hold no golden numbers
(pulsemodel/test/test_smoke.py:139-214 only checks run-to-run repeatability and the (-1,65)/(-1,33) shapes).
This file therefore *defines* the operator from what the call sites fix:

  call sites   percivaltts/vocoders.py:134-145 (compress_spectrum: log|SPEC| -> linbnd2fwbnd),
               percivaltts/vocoders.py:147-166 (decompress_spectrum: exp(fwbnd2linbnd(., smooth=True))),
               percivaltts/vocoders.py:202 (NM: fwbnd2linbnd, no smoothing),
               percivaltts/external/pulsemodel/analysis.py:336,366, synthesis.py:316,351,
               percivaltts/optimizertts_wgan.py:192,200 (freq2fwspecidx)
  constraints  "mel-bands" (analysis.py:284,287); exactly nbbnds columns including the DC and Nyquist ends
               (odd sizes 129/65/33/17, analysis.py:388); the noise mask stays in [0,1] through compress and
               decompress (synthesis.py:82,376) => a convex, interpolating operator; a `smooth=True` variant
               exists for log-spectra; dftlen = (cols-1)*2 (vocoders.py:136).

Definition (all tables built in float64):
  mel(f)   = 1127 ln(1 + f/700)
  band k   sits at  f_k = mel^-1( k/(K-1) * mel(fs/2) ),  k = 0..K-1      (K = nbbnds)
  linbnd2fwbnd(X)[t,k]      = linear interpolation of the row X[t,:] (bins at i*fs/dftlen) at f_k
  fwbnd2linbnd(Y)[t,i]      = linear interpolation of the row Y[t,:] (bands at u=0..K-1) at
                              u_i = (K-1) * mel(i*fs/dftlen) / mel(fs/2)
  fwbnd2linbnd(Y, smooth)   = Catmull-Rom cubic through the same uniform band grid (end bands replicated)
  freq2fwspecidx(f, fs, K)  = (K-1) * mel(f)/mel(fs/2)   (fractional band position)
The product exposes the tables (index + weight per output column) as kernel inputs, so a user holding the real
sigproc can substitute its tables without touching the kernels.
"""
import numpy as np


def mel(f):
    return 1127.0 * np.log1p(np.asarray(f, dtype=np.float64) / 700.0)


def imel(m):
    return 700.0 * np.expm1(np.asarray(m, dtype=np.float64) / 1127.0)


def freq2fwspecidx(freq, fs, nbbnds):
    return (nbbnds - 1) * mel(freq) / mel(0.5 * fs)


def band_positions(fs, dftlen, nbbnds):
    """Fractional linear-bin position of every band centre."""
    k = np.arange(nbbnds, dtype=np.float64)
    fk = imel(k / (nbbnds - 1) * mel(0.5 * fs))
    p = fk * dftlen / float(fs)
    p[0] = 0.0
    p[-1] = dftlen // 2
    return p


def bin_positions(fs, dftlen, nbbnds):
    """Fractional band position of every linear bin."""
    i = np.arange(dftlen // 2 + 1, dtype=np.float64)
    u = (nbbnds - 1) * mel(i * fs / float(dftlen)) / mel(0.5 * fs)
    u[0] = 0.0
    u[-1] = nbbnds - 1
    return u


def linbnd2fwbnd(X, fs, dftlen, nbbnds):
    X = np.asarray(X, dtype=np.float64)
    nbins = dftlen // 2 + 1
    assert X.shape[1] == nbins
    p = band_positions(fs, dftlen, nbbnds)
    bins = np.arange(nbins, dtype=np.float64)
    out = np.empty((X.shape[0], nbbnds), dtype=np.float64)
    for t in range(X.shape[0]):
        out[t] = np.interp(p, bins, X[t])
    return out


def fwbnd2linbnd(Y, fs, dftlen, smooth=False):
    Y = np.asarray(Y, dtype=np.float64)
    K = Y.shape[1]
    u = bin_positions(fs, dftlen, K)
    out = np.empty((Y.shape[0], dftlen // 2 + 1), dtype=np.float64)
    if not smooth:
        bands = np.arange(K, dtype=np.float64)
        for t in range(Y.shape[0]):
            out[t] = np.interp(u, bands, Y[t])
        return out
    j = np.minimum(np.floor(u).astype(np.int64), K - 2)
    w = u - j
    jm, j1, j2 = np.maximum(j - 1, 0), j + 1, np.minimum(j + 2, K - 1)
    w2, w3 = w * w, w * w * w
    for t in range(Y.shape[0]):
        p0, p1, p2, p3 = Y[t, jm], Y[t, j], Y[t, j1], Y[t, j2]
        out[t] = 0.5 * (2.0 * p1 + (p2 - p0) * w + (2.0 * p0 - 5.0 * p1 + 4.0 * p2 - p3) * w2
                        + (3.0 * p1 - p0 - 3.0 * p2 + p3) * w3)
    return out


def compress_spectrum(SPEC, fs, spec_size):
    """vocoders.py:134-145 ('fwbnd' branch)"""
    dftlen = (SPEC.shape[1] - 1) * 2
    return linbnd2fwbnd(np.log(np.abs(np.asarray(SPEC, dtype=np.float64))), fs, dftlen, spec_size)


def decompress_spectrum(COMPSPEC, fs, dftlen):
    """vocoders.py:147-150 ('fwbnd' branch, pp_mcep=False)"""
    return np.exp(fwbnd2linbnd(COMPSPEC, fs, dftlen, smooth=True))


def generation_tail(CMP, mean, std, fs, dftlen, spec_size, nm_size):
    """De-normalisation + vocoder front half of generation, float64:
    modeltts.py:156-158 (CMP*std + mean), vocoders.py:194-202 (f0 = exp(CMP[:,0]); SPEC = decompress_spectrum;
    NM = fwbnd2linbnd(CMP[:, 1+S:1+S+NM])).  -> (CMP de-normalised, f0, SPEC, NM)"""
    CMP = np.asarray(CMP, dtype=np.float64)
    if mean is not None:
        CMP = CMP * np.asarray(std, dtype=np.float64)[np.newaxis] + np.asarray(mean, dtype=np.float64)[np.newaxis]
    f0 = np.exp(CMP[:, 0])
    SPEC = decompress_spectrum(CMP[:, 1:1 + spec_size], fs, dftlen)
    NM = fwbnd2linbnd(CMP[:, 1 + spec_size:1 + spec_size + nm_size], fs, dftlen)
    return CMP, f0, SPEC, NM
