"""
Vocoder descriptors on the hot path: feature-vector sizes and the PML spectral-envelope / noise-mask band
compression.  Mirrors reference percivaltts/vocoders.py:30-218 (Vocoder, VocoderF0Spec, VocoderPML) and the size API
of VocoderWORLD (:221-343).  Waveform analysis / synthesis (REAPER, WORLD, pulsemodel) are out of scope: the methods
exist and raise NotImplementedError.
"""
import numpy as np

from . import sigproc as sp


class Vocoder(object):
    def __init__(self, name, fs, shift, mlpg_wins=None):
        self._name = name
        self.fs = fs
        self.shift = shift
        self.mlpg_wins = mlpg_wins
        self.features_err = dict()

    def __str__(self):
        return '{} (fs={}, shift={})'.format(self.name(), self.fs, self.shift)

    def name(self):
        return self._name

    def featuressizeraw(self):
        raise ValueError('This member function has to be re-implemented in the sub-classes')

    def featuressize(self):
        """vocoders.py:102-104: static features, times (1 + number of MLPG windows)."""
        if self.mlpg_wins is not None:
            return self.featuressizeraw() * (len(self.mlpg_wins) + 1)
        return self.featuressizeraw()

    def f0size(self):
        return -1

    def specsize(self):
        return -1

    def noisesize(self):
        return -1

    def vuvsize(self):
        return -1

    def objmeasures_clear(self):
        self.features_err = dict()

    def objmeasures_stats(self):
        for key in self.features_err:
            print('{}: {}'.format(key, np.mean(np.vstack(self.features_err[key]))))


class VocoderF0Spec(Vocoder):
    def __init__(self, name, fs, shift, spec_size, spec_type='fwbnd', dftlen=4096, mlpg_wins=None):
        Vocoder.__init__(self, name, fs, shift, mlpg_wins=mlpg_wins)
        self.spec_size = spec_size
        self.spec_type = spec_type
        self.dftlen = dftlen

    def f0size(self):
        return 1

    def specsize(self):
        return self.spec_size

    def compress_spectrum(self, SPEC, spec_type=None, spec_size=None):
        """[T, dftlen/2+1] amplitude spectra -> [T, spec_size] log-amplitude warped bands (vocoders.py:134-145).
        The log|.| is fused into the band kernel."""
        spec_size = self.spec_size if spec_size is None else spec_size
        if self.spec_type != 'fwbnd':
            raise NotImplementedError('only spec_type="fwbnd" is on the B200 path (mcep needs SPTK)')
        dftlen = (SPEC.shape[1] - 1) * 2
        ops = sp.backend_b200.get_ops()
        return sp.linbnd2fwbnd_device(sp._to_device(np.abs(SPEC), ops), self.fs, dftlen, spec_size, log_abs=True,
                                      ops=ops).cpu().numpy()

    def decompress_spectrum(self, COMPSPEC, spec_type=None, pp_mcep=False):
        """[T, spec_size] -> [T, dftlen/2+1] amplitude spectra: exp(fwbnd2linbnd(., smooth=True)) (vocoders.py:147-150)."""
        if self.spec_type != 'fwbnd' or pp_mcep:
            raise NotImplementedError('only spec_type="fwbnd" without MCEP post-processing is on the B200 path')
        ops = sp.backend_b200.get_ops()
        return sp.fwbnd2linbnd_device(sp._to_device(COMPSPEC, ops), self.fs, self.dftlen, smooth=True, do_exp=True,
                                      ops=ops).cpu().numpy()


class VocoderPML(VocoderF0Spec):
    def __init__(self, fs, shift, spec_size, nm_size, dftlen=4096, mlpg_wins=None):
        VocoderF0Spec.__init__(self, 'PML', fs, shift, spec_size, 'fwbnd', dftlen, mlpg_wins=mlpg_wins)
        self.nm_size = nm_size

    def featuressizeraw(self):
        return 1 + self.spec_size + self.nm_size

    def noisesize(self):
        return self.nm_size

    def compress_noisemask(self, NM, nm_size=None):
        """[T, dftlen/2+1] noise mask in [0,1] -> [T, nm_size] (pulsemodel/analysis.py:366)."""
        dftlen = (NM.shape[1] - 1) * 2
        return sp.linbnd2fwbnd(NM, self.fs, dftlen, self.nm_size if nm_size is None else nm_size)

    def decompress_noisemask(self, NM):
        """[T, nm_size] -> [T, dftlen/2+1], plain linear interpolation (vocoders.py:202)."""
        return sp.fwbnd2linbnd(NM, self.fs, self.dftlen)

    def decompress_features_device(self, cmp, mean=None, std=None, ops=None, keep_cmp=False):
        """The generation tail on the device, chained without leaving HBM:
            cmp [T,O] (device, the generator's output) -> de-normalise with mean4norm / std4norm (modeltts.py:154-165) ->
            f0 = exp(.[:,0]), SPEC = exp(fwbnd2linbnd(.[:,1:1+S], smooth)), NM = fwbnd2linbnd(.[:,1+S:]) (vocoders.py:194-202).
        Returns one flat device buffer laid out [f0 (T) | SPEC (T*bins) | NM (T*bins) | de-normalised cmp (T*O) if keep_cmp] and
        the views (f0, SPEC, NM, CMP or None) into it: ONE device->host copy brings everything the synthesiser needs."""
        import torch
        ops = ops or sp.backend_b200.get_ops()
        T, O = cmp.shape
        S, NMs, bins = self.spec_size, self.nm_size, self.dftlen // 2 + 1
        if 1 + S + NMs > O:
            raise ValueError('feature vectors of size {} cannot hold f0 + {} + {} bands'.format(O, S, NMs))
        flat = ops.empty(T * (1 + 2 * bins + (O if keep_cmp else 0)))
        f0 = flat[:T]
        SPEC = flat[T:T + T * bins].view(T, bins)
        NM = flat[T + T * bins:T + 2 * T * bins].view(T, bins)
        den = flat[T + 2 * T * bins:].view(T, O) if keep_cmp else None
        if T == 0:
            return flat, (f0, SPEC, NM, den)
        spec_c, nm_c = ops.empty(T, S), ops.empty(T, NMs)
        md = sd = None
        if mean is not None:
            md = torch.as_tensor(np.asarray(mean, dtype=np.float32)).to(ops.device) if not torch.is_tensor(mean) else mean
            sd = torch.as_tensor(np.asarray(std, dtype=np.float32)).to(ops.device) if not torch.is_tensor(std) else std
            if md.numel() != O or sd.numel() != O:
                raise ValueError('mean4norm / std4norm have {} / {} entries for {} features'.format(md.numel(), sd.numel(), O))
        ops.denorm_split(cmp, md, sd, S, NMs, out=den, f0=f0, spec=spec_c, nm=nm_c)
        sp.fwbnd2linbnd_device(spec_c, self.fs, self.dftlen, smooth=True, do_exp=True, out=SPEC, ops=ops)
        sp.fwbnd2linbnd_device(nm_c, self.fs, self.dftlen, out=NM, ops=ops)
        return flat, (f0, SPEC, NM, den)

    def decompress_features(self, CMP, mean=None, std=None):
        """The device-side part of synthesis() (vocoders.py:194-202): -> (f0 [T], SPEC [T,bins], NM [T,bins]) numpy.
        One host->device copy of the [T,O] features, the chain above, one device->host copy."""
        ops = sp.backend_b200.get_ops()
        CMP = np.ascontiguousarray(CMP, dtype=np.float32)
        T, bins = CMP.shape[0], self.dftlen // 2 + 1
        flat, _ = self.decompress_features_device(sp._to_device(CMP, ops), mean, std, ops=ops)
        host = flat.cpu().numpy()
        return host[:T], host[T:T + T * bins].reshape(T, bins), host[T + T * bins:T + 2 * T * bins].reshape(T, bins)

    def objmeasures_add(self, CMP, REF):
        """vocoders.py:208-218: RMS errors of f0 [Hz], spectral bands [dB] and noise mask between generated and reference features."""
        log2db = 20.0 / np.log(10.0)
        S, NMs = self.spec_size, self.nm_size
        self.features_err.setdefault('F0[Hz]', []).append(np.sqrt(np.mean((np.exp(REF[:, 0]) - np.exp(CMP[:, 0])) ** 2)))
        self.features_err.setdefault('SPEC[dB]', []).append(
            np.sqrt(np.mean((log2db * REF[:, 1:1 + S] - log2db * CMP[:, 1:1 + S]) ** 2, 0)))
        self.features_err.setdefault('NM', []).append(np.sqrt(np.mean((REF[:, 1 + S:1 + S + NMs] - CMP[:, 1 + S:1 + S + NMs]) ** 2, 0)))

    def analysisf(self, *args, **kwargs):
        raise NotImplementedError('waveform analysis (REAPER / pulsemodel) is outside the B200 hot path')

    def synthesis(self, CMP, pp_mcep=False, pp_f0_smooth=None, synthesizer=None):
        """vocoders.py:194-206.  Everything up to pulsemodel.synthesis.synthesize runs on the device (decompress_features);
        the pulse-by-pulse waveform synthesis itself is outside the B200 hot path: pass it in as
        `synthesizer(fs, f0s [T,2] = (time, f0), SPEC, NM) -> waveform`."""
        if pp_mcep:
            raise NotImplementedError('MCEP post-processing (SPTK) is not on the B200 path')
        if synthesizer is None:
            raise NotImplementedError('pulse-by-pulse waveform synthesis (pulsemodel) is outside the B200 hot path; pass '
                                      'synthesizer=, or use decompress_features() for its inputs')
        f0, SPEC, NM = self.decompress_features(CMP)
        return synthesizer(self.fs, np.vstack((self.shift * np.arange(len(f0)), f0)).T, SPEC, NM)


class VocoderWORLD(VocoderF0Spec):
    """Size API only (vocoders.py:221-240): f0 + spec + aperiodicity bands + vuv."""

    def __init__(self, fs, shift, spec_size, aper_size, dftlen=4096, mlpg_wins=None):
        VocoderF0Spec.__init__(self, 'WORLD', fs, shift, spec_size, 'fwbnd', dftlen, mlpg_wins=mlpg_wins)
        self.aper_size = aper_size

    def featuressizeraw(self):
        return 1 + self.spec_size + self.aper_size + 1

    def noisesize(self):
        return self.aper_size

    def vuvsize(self):
        return 1
