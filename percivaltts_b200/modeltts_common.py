"""
Model definitions (reference percivaltts/modeltts_common.py:36-126).

DCNNF0SpecNoiseFeatures is the paper's generator; its graph is executed by engine.Generator.
Generic (FC layers) is executed by engine.GenericFC; with OptimizerTTS's LSE hooks it is the reference's plain
least-squares training path (optimizertts.py:378-436).  count_params()==2195 is the reference's pinned known answer.
"""
from . import arch as A
from . import modeltts


class Generic(modeltts.ModelTTS):
    def __init__(self, ctxsize, vocoder, fmodel=None, layertypes=['FC', 'FC', 'FC'], nameprefix=None, cfgarch=None):
        modeltts.ModelTTS.__init__(self, ctxsize, vocoder)
        if fmodel is not None:
            self.load(fmodel)
            self.kerasmodel.summary()
            return
        if vocoder.mlpg_wins is not None:
            raise NotImplementedError('MLPG output layers are not on the B200 path')
        arch = A.GenericArch(ctxsize, vocoder.specsize(), vocoder.noisesize(), cfgarch.arch_hiddenwidth, layertypes)
        shapes = A.generic_param_shapes(ctxsize, vocoder.specsize(), vocoder.noisesize(), cfgarch.arch_hiddenwidth,
                                        layertypes)
        self.kerasmodel = modeltts.NativeNetwork('generic', arch, shapes)
        self.kerasmodel.summary()


class DCNNF0SpecNoiseFeatures(modeltts.ModelTTS):
    """The model of the PercivalTTS SLT paper (modeltts_common.py:61-126)."""

    def __init__(self, ctxsize, vocoder, cfgarch, nameprefix=None):
        modeltts.ModelTTS.__init__(self, ctxsize, vocoder)
        arch = A.Arch.from_cfg(ctxsize, vocoder, cfgarch)
        self.kerasmodel = modeltts.NativeNetwork('generator', arch, A.generator_param_shapes(arch))
        self.kerasmodel.summary()
