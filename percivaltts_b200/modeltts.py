"""
Model wrapper: the python-3 / B200 counterpart of reference percivaltts/modeltts.py:46-205 (ModelTTS).

`kerasmodel` keeps its name (the optimizer touches it, optimizertts_wgan.py:111,121,316,328) but holds a
NativeNetwork: parameter inventory + host weights, materialised on the GPU (engine.Generator / engine.Critic on
libpctts.so) the first time it is used.  Construction and count_params() need no GPU.
"""
import json
import os
import pickle
import sys
from collections import OrderedDict

import numpy as np
import torch

from . import arch as A
from . import backend_b200


class NativeNetwork(object):
    """Stands where the reference has a keras.Model."""

    def __init__(self, kind, arch, shapes, seed=None):
        self.kind = kind                  # 'generator' | 'critic' | 'generic'
        self.arch = arch
        self.shapes = list(shapes)
        self._host = A.init_params(self.shapes, backend_b200.get_seed() if seed is None else seed)
        self._module = None
        self._predict_shapes = OrderedDict()      # LRU of (B, T) inference plans (one T per utterance at validation / generation)
        self.max_predict_shapes = 16

    # -- keras.Model-like surface -------------------------------------------------------------------------------
    def count_params(self):
        return A.count_params(self.shapes)

    def summary(self):
        print('    {} network: {} parameter tensors, {} parameters'.format(self.kind, len(self.shapes), self.count_params()))
        for name, shape, trainable in self.shapes:
            print('        {:<34}{:<22}{}'.format(name, str(tuple(shape)), '' if trainable else '(non-trainable)'))

    def get_weights(self):
        """OrderedDict name -> float32 numpy array (read back from the device when materialised)."""
        if self._module is not None:
            return self._module.params.export()
        return OrderedDict((k, v.copy()) for k, v in self._host.items())

    def set_weights(self, weights):
        for name, shape, _ in self.shapes:
            w = np.asarray(weights[name], dtype=np.float32)
            if tuple(w.shape) != tuple(shape):
                raise ValueError('shape mismatch for {}: {} vs {}'.format(name, w.shape, shape))
            self._host[name] = w.copy()
        if self._module is not None:
            self._module.params.load(self._host)

    def device_module(self, ops=None):
        """engine.Generator / engine.Critic holding the parameters on the GPU (created once)."""
        if self._module is None:
            from . import engine
            ops = ops or backend_b200.get_ops()
            if self.kind == 'generator':
                self._module = engine.Generator(ops, self.arch)
            elif self.kind == 'critic':
                self._module = engine.Critic(ops, self.arch)
            elif self.kind == 'generic':
                self._module = engine.GenericFC(ops, self.arch)
            else:
                raise ValueError('unknown network kind ' + str(self.kind))
            self._module.params.load(self._host)
        return self._module

    def predict_device(self, x):
        """numpy or device [B,T,ctxsize] -> DEVICE tensor [B,T,out], BatchNorm in inference mode (moving statistics)."""
        if self.kind not in ('generator', 'generic'):
            raise NotImplementedError('predict() is defined for the generator')
        G = self.device_module()
        if not torch.is_tensor(x):
            x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 3 or x.shape[2] != G.a.ctxsize:
            raise ValueError('expected input of shape [batch, frames, {}]'.format(G.a.ctxsize))
        B, T, _ = x.shape
        out = G.ops.empty(B, T, G.a.outsize)
        if T == 0 or B == 0:
            return out
        xd = x.to(G.ops.device).contiguous() if torch.is_tensor(x) else torch.from_numpy(x).to(G.ops.device)
        tag = ('predict', B, T)
        self._predict_shapes[tag] = True
        self._predict_shapes.move_to_end(tag)
        while len(self._predict_shapes) > self.max_predict_shapes:        # inference plans + their kernel scratch, oldest first
            old, _ = self._predict_shapes.popitem(last=False)
            G.release(old[1], old[2], with_backward=False)
            G.ops.release_shape(old)
        with G.ops.shape_scope(tag), G.ops.step_scope():
            G.forward(G.plan(B, T, with_backward=False), xd, out, training=False)
        return out

    def predict(self, x):
        """numpy [B,T,ctxsize] -> numpy [B,T,out], BatchNorm in inference mode (moving statistics)."""
        return self.predict_device(x).cpu().numpy()

    def to_json(self):
        return json.dumps(dict(kind=self.kind, arch=self.arch.as_dict() if self.arch is not None else None,
                               shapes=[[n, list(s), bool(t)] for n, s, t in self.shapes]), indent=1)

    def save_weights(self, fname):
        np.savez(fname, **self.get_weights())

    def load_weights(self, fname):
        with np.load(fname if fname.endswith('.npz') else fname + '.npz') as z:
            self.set_weights(dict((k, z[k]) for k in z.files))


def network_from_json(txt):
    d = json.loads(txt)
    arch = None
    if d.get('arch'):
        arch = A.GenericArch(**d['arch']) if d['kind'] == 'generic' else A.Arch(**d['arch'])
    return NativeNetwork(d['kind'], arch, [(n, tuple(s), bool(t)) for n, s, t in d['shapes']])


class ModelTTS(object):
    ctxsize = -1
    vocoder = None
    kerasmodel = None

    def __init__(self, ctxsize, vocoder, kerasmodel=None):
        print('Building the TTS-dedicated model')
        self.ctxsize = ctxsize
        self.vocoder = vocoder
        if kerasmodel is not None:
            self.kerasmodel = kerasmodel
            self.kerasmodel.summary()

    def predict(self, x):
        return self.kerasmodel.predict(x)                                 # modeltts.py:68-69

    def count_params(self):
        return self.kerasmodel.count_params()                             # modeltts.py:72-73

    def save(self, fmodel, cfg=None, extras=None, printfn=print, infostr=''):
        """Same three files as modeltts.py:75-105, with a numpy container instead of HDF5:
        <fmodel>.arch.json, <fmodel>.weights.npz, <fmodel>.cfgextras.pkl"""
        if extras is None:
            extras = dict()
        printfn('    saving parameters in {} ...'.format(fmodel), end='')
        with open(fmodel + '.arch.json', 'w') as f:
            f.write(self.kerasmodel.to_json())
        self.kerasmodel.save_weights(fmodel + '.weights.npz')
        with open(fmodel + '.cfgextras.pkl', 'wb') as f:
            pickle.dump([cfg, extras], f)
        print(' done ' + infostr)
        sys.stdout.flush()

    def load(self, fmodel, printfn=print, compile=True):
        printfn('    reloading parameters from {} ...'.format(fmodel), end='')
        if self.kerasmodel is None:
            with open(fmodel + '.arch.json') as f:
                self.kerasmodel = network_from_json(f.read())
        self.kerasmodel.load_weights(fmodel + '.weights.npz')
        with open(fmodel + '.cfgextras.pkl', 'rb') as f:
            DATA = pickle.load(f)
        print(' done')
        sys.stdout.flush()
        return DATA

    def generate_cmp(self, inpath, outpath, fid_lst):
        """One utterance at a time, raw float32 [T, out] files (modeltts.py:133-141)."""
        from . import data
        outdir = os.path.dirname(outpath)
        if outdir and not os.path.isdir(outdir):
            os.makedirs(outdir)
        X = data.load(inpath, fid_lst, verbose=1)
        for vi, fid in enumerate(fid_lst):
            CMP = self.predict(X[vi][np.newaxis])
            CMP[0].astype('float32').tofile(outpath.replace('*', fid))

    def generate_features_device(self, x, mean=None, std=None, keep_cmp=True):
        """One utterance through the whole generation tail on the device: x numpy [T,ctxsize] -> predict (BN inference mode) ->
        de-normalise (mean4norm / std4norm, reference modeltts.py:154-165) -> exp(f0) / decompress_spectrum / NM fwbnd2linbnd
        (vocoders.py:194-202), one H2D of the labels and one D2H of the result.  -> (CMP [T,O] de-normalised, f0 [T],
        SPEC [T,bins], NM [T,bins]) numpy views of one host buffer."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2:
            raise ValueError('expected one utterance [frames, ctxsize]')
        T, O = x.shape[0], self.vocoder.featuressize()
        bins = self.vocoder.dftlen // 2 + 1
        pred = self.kerasmodel.predict_device(x[np.newaxis])            # device [1,T,O]
        flat, _ = self.vocoder.decompress_features_device(pred.view(T, O), mean, std, ops=self.kerasmodel.device_module().ops,
                                                          keep_cmp=keep_cmp)
        host = flat.cpu().numpy()
        f0, SPEC = host[:T], host[T:T + T * bins].reshape(T, bins)
        NM = host[T + T * bins:T + 2 * T * bins].reshape(T, bins)
        CMP = host[T + 2 * T * bins:].reshape(T, O) if keep_cmp else None
        return CMP, f0, SPEC, NM

    def generate_wav(self, inpath, outpath, fid_lst, syndir, do_objmeas=True, do_resynth=True, pp_mcep=False,
                     pp_spec_pf_coef=-1, pp_spec_extrapfreq=-1, pp_f0_smooth=None, synthesizer=None, wavwrite=None):
        """Generation loop of reference modeltts.py:144-205 with the tail on the device: per utterance predict -> de-normalise with
        <dirname(outpath)>/mean4norm.dat, std4norm.dat -> vocoder inputs (f0, SPEC, NM), then the waveform synthesiser.
        Pulse-by-pulse synthesis (pulsemodel) is outside the B200 path: `synthesizer(fs, f0s, SPEC, NM) -> waveform` and
        `wavwrite(path, wav, fs)` are injected; without a synthesizer the vocoder inputs are written as raw float32 files
        <syndir>/<fid>.{cmp,f0,spec,nm} (the on-wire format of README.md:184-188) for an external pulsemodel run."""
        from . import data
        if pp_mcep or pp_spec_pf_coef > 0 or pp_spec_extrapfreq > 0:
            raise NotImplementedError('spectral post-processing (SPTK / formant enhancement) is not on the B200 path')
        if self.vocoder.mlpg_wins is not None and len(self.vocoder.mlpg_wins) > 0:
            raise NotImplementedError('MLPG (merlin) is not on the B200 path')
        print('Reloading output stats')
        Ymean = np.fromfile(os.path.join(os.path.dirname(outpath), 'mean4norm.dat'), dtype='float32')
        Ystd = np.fromfile(os.path.join(os.path.dirname(outpath), 'std4norm.dat'), dtype='float32')
        print('\nLoading generation data at once ...')
        X_test = data.load(inpath, fid_lst, verbose=1)
        y_test = None
        if do_objmeas or do_resynth:
            y_test = data.load(outpath, fid_lst, verbose=1)
            X_test, y_test = data.croplen((X_test, y_test))
        if not os.path.isdir(syndir):
            os.makedirs(syndir)
        if do_resynth and synthesizer is not None and not os.path.isdir(syndir + '-resynth'):
            os.makedirs(syndir + '-resynth')
        self.vocoder.objmeasures_clear()
        for vi, fid in enumerate(fid_lst):
            print('Generating {}/{} fid={} ...'.format(1 + vi, len(X_test), fid))
            if do_resynth and synthesizer is not None:
                f0, SPEC, NM = self.vocoder.decompress_features(y_test[vi], Ymean, Ystd)
                resyn = synthesizer(self.vocoder.fs, np.vstack((self.vocoder.shift * np.arange(len(f0)), f0)).T, SPEC, NM)
                if wavwrite is not None:
                    wavwrite(os.path.join(syndir + '-resynth', fid + '.wav'), resyn, self.vocoder.fs)
            CMP, f0, SPEC, NM = self.generate_features_device(X_test[vi], Ymean, Ystd)
            if synthesizer is not None:
                syn = synthesizer(self.vocoder.fs, np.vstack((self.vocoder.shift * np.arange(len(f0)), f0)).T, SPEC, NM)
                if wavwrite is not None:
                    wavwrite(os.path.join(syndir, fid + '.wav'), syn, self.vocoder.fs)
            else:
                for ext, arr in (('cmp', CMP), ('f0', f0), ('spec', SPEC), ('nm', NM)):
                    np.ascontiguousarray(arr, dtype=np.float32).tofile(os.path.join(syndir, fid + '.' + ext))
            if do_objmeas:
                ref = y_test[vi] * Ystd + Ymean
                self.vocoder.objmeasures_add(CMP, ref)
        if do_objmeas:
            self.vocoder.objmeasures_stats()
        print('Generation finished')
