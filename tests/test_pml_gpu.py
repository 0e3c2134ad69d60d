"""
GPU parity of the PML band compression / decompression kernels (csrc/pml.cu, through the C ABI) against the CPU oracle
(oracle/pml_oracle.py) and the committed golden fixture tests/golden/pml_golden.npz.

Tolerance (north_star): 1e-5.  Band-domain outputs and noise masks: max |gpu - oracle| <= 1e-5 * max |oracle| per call.
Decompressed amplitude spectra (exp of the interpolated log-amplitudes, 8 decades of dynamic range): ELEMENTWISE relative,
|gpu - oracle| <= 1e-5 * |oracle| for every bin.
Reference call sites: percivaltts/vocoders.py:134-166,202; pulsemodel/analysis.py:336,366.
"""
import os

import numpy as np
import pytest
import torch

from oracle import pml_oracle as P
from percivaltts_b200 import sigproc as sp
from percivaltts_b200 import vocoders

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
FS, DFTLEN, NB = 16000, 4096, 2049
TOL = 1e-5


def rel(got, want):
    want = np.asarray(want, dtype=np.float64)
    return float(np.abs(np.asarray(got, dtype=np.float64) - want).max() / (np.abs(want).max() + 1e-300))


def relel(got, want):
    want = np.asarray(want, dtype=np.float64)
    return float((np.abs(np.asarray(got, dtype=np.float64) - want) / np.abs(want)).max())


def synth(T, seed):
    rng = np.random.RandomState(seed)
    spec = np.exp(rng.normal(-6.0, 2.0, size=(T, NB))).astype(np.float32)          # BASELINE config #5
    nm = rng.uniform(0, 1, size=(T, NB))
    nm = (np.convolve(nm.ravel(), np.ones(9) / 9.0, mode='same').reshape(T, NB) > 0.5).astype(np.float32)
    return spec, nm


def test_golden_fixture(cuda_ops):
    z = np.load(os.path.join(GOLD, 'pml_golden.npz'))
    for K in (129, 65):
        voc = vocoders.VocoderPML(FS, 0.005, K, 33)
        c = voc.compress_spectrum(z['SPEC'])
        assert c.shape == (6, K) and c.dtype == np.float32 and rel(c, z['cspec%d' % K]) < TOL
        d = voc.decompress_spectrum(z['cspec%d' % K].astype(np.float32))
        assert d.shape == (6, NB) and relel(d, z['dspec%d' % K]) < TOL
    for K in (33, 17):
        voc = vocoders.VocoderPML(FS, 0.005, 129, K)
        c = voc.compress_noisemask(z['NM'])
        d = voc.decompress_noisemask(z['cnm%d' % K].astype(np.float32))
        assert rel(c, z['cnm%d' % K]) < TOL and rel(d, z['dnm%d' % K]) < TOL
        assert c.min() >= 0.0 and c.max() <= 1.0 and d.min() >= 0.0 and d.max() <= 1.0        # synthesis.py:82,376


@pytest.mark.parametrize('T', [1, 3, 15, 16, 17, 33, 257, 1000])     # ragged against the kernels' 4-warp / 16-frame tiles
def test_against_oracle_ragged(cuda_ops, T):
    spec, nm = synth(T, T)
    for K in (129, 65):
        c = sp.linbnd2fwbnd_device(torch.from_numpy(spec).cuda(), FS, DFTLEN, K, log_abs=True).cpu().numpy()
        want = P.compress_spectrum(spec, FS, K)
        assert rel(c, want) < TOL
        d = sp.fwbnd2linbnd_device(torch.from_numpy(want.astype(np.float32)).cuda(), FS, DFTLEN, smooth=True, do_exp=True)
        wantd = P.decompress_spectrum(want.astype(np.float32), FS, DFTLEN)
        assert relel(d.cpu().numpy(), wantd) < TOL
        # smooth without exp, and plain linear with exp: the two remaining template instances
        d2 = sp.fwbnd2linbnd_device(torch.from_numpy(want.astype(np.float32)).cuda(), FS, DFTLEN, smooth=True, do_exp=False)
        assert rel(d2.cpu().numpy(), P.fwbnd2linbnd(want.astype(np.float32), FS, DFTLEN, smooth=True)) < TOL
        d3 = sp.fwbnd2linbnd_device(torch.from_numpy(want.astype(np.float32)).cuda(), FS, DFTLEN, smooth=False, do_exp=True)
        assert relel(d3.cpu().numpy(), np.exp(P.fwbnd2linbnd(want.astype(np.float32), FS, DFTLEN))) < TOL
    for K in (33, 17):
        c = sp.linbnd2fwbnd(nm, FS, DFTLEN, K)
        want = P.linbnd2fwbnd(nm, FS, DFTLEN, K)
        assert rel(c, want) < TOL
        d = sp.fwbnd2linbnd(want.astype(np.float32), FS, DFTLEN)
        assert rel(d, P.fwbnd2linbnd(want.astype(np.float32), FS, DFTLEN)) < TOL
        assert d.min() >= 0.0 and d.max() <= 1.0


def test_empty_and_other_dft_sizes(cuda_ops):
    out = sp.linbnd2fwbnd(np.zeros((0, NB), dtype=np.float32), FS, DFTLEN, 129)
    assert out.shape == (0, 129)
    out = sp.fwbnd2linbnd(np.zeros((0, 33), dtype=np.float32), FS, DFTLEN)
    assert out.shape == (0, NB)
    rng = np.random.RandomState(3)
    for dftlen, K in ((512, 17), (1024, 65), (8192, 129)):
        x = rng.uniform(0.1, 2.0, size=(37, dftlen // 2 + 1)).astype(np.float32)
        c = sp.linbnd2fwbnd(x, 44100, dftlen, K)
        want = P.linbnd2fwbnd(x, 44100, dftlen, K)
        assert rel(c, want) < TOL
        assert rel(sp.fwbnd2linbnd(want.astype(np.float32), 44100, dftlen, smooth=True),
                   P.fwbnd2linbnd(want.astype(np.float32), 44100, dftlen, smooth=True)) < TOL


def test_full_size_properties(cuda_ops):
    """BASELINE config #5 scale (a 64 k-frame slab of the 1 M frames; the oracle cannot follow there): constants are fixed
    points, the noise mask stays in [0,1], the DC / Nyquist bands are the DC / Nyquist bins, the operator is linear."""
    T = 65536
    g = torch.Generator(device='cuda').manual_seed(11)
    nm = torch.rand(T, NB, device='cuda', generator=g)
    c = sp.linbnd2fwbnd_device(nm, FS, DFTLEN, 33)
    d = sp.fwbnd2linbnd_device(c, FS, DFTLEN)
    assert float(c.min()) >= 0.0 and float(c.max()) <= 1.0 and float(d.min()) >= 0.0 and float(d.max()) <= 1.0
    assert torch.equal(c[:, 0], nm[:, 0]) and torch.equal(c[:, -1], nm[:, -1])
    assert torch.equal(d[:, 0], c[:, 0]) and float((d[:, -1] - c[:, -1]).abs().max()) <= 1.2e-7
    const = torch.full((T, NB), 0.37, device='cuda')
    cc = sp.linbnd2fwbnd_device(const, FS, DFTLEN, 129)
    assert float((cc - 0.37).abs().max()) < 1e-6
    assert float((sp.fwbnd2linbnd_device(cc, FS, DFTLEN, smooth=True) - 0.37).abs().max()) < 1e-6
    a, b = torch.rand(T, 129, device='cuda', generator=g), torch.rand(T, 129, device='cuda', generator=g)
    lhs = sp.fwbnd2linbnd_device(a + b, FS, DFTLEN, smooth=True)
    rhs = sp.fwbnd2linbnd_device(a, FS, DFTLEN, smooth=True) + sp.fwbnd2linbnd_device(b, FS, DFTLEN, smooth=True)
    assert float((lhs - rhs).abs().max()) < 5e-6
    # every frame of the slab was written (ragged tail included): compare the last tile with the same rows done alone
    tail = sp.fwbnd2linbnd_device(a[-5:].contiguous(), FS, DFTLEN, smooth=True)
    assert torch.equal(tail, sp.fwbnd2linbnd_device(a, FS, DFTLEN, smooth=True)[-5:])
