"""
Writes tests/golden/data_golden.npz: the outputs of the REFERENCE's own percivaltts/data.py and percivaltts/percivaltts.py
(/root/reference, unmodified, executed in this container) on the seeded scenario of tests/golden/data_scenarios.py.
    python tests/golden/make_data_golden.py

The reference modules are python-2 era numpy code: they import under python 3 once `xrange` exists and the two optional
imports of percivaltts.py that this image lacks (defusedxml, matplotlib -- used by its log plots only) are stubbed.  Nothing of
the reference is copied: only its results are stored.
"""
import builtins
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = '/root/reference/percivaltts'


def import_reference():
    builtins.xrange = range
    for name in ('defusedxml', 'defusedxml.ElementTree', 'matplotlib', 'matplotlib.pyplot'):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    if hasattr(sys.modules.get('matplotlib'), 'use') is False:
        sys.modules['matplotlib'].use = lambda *a, **k: None
    sys.path.insert(0, REF)
    import percivaltts as pt      # /root/reference/percivaltts/percivaltts.py (the scripts run from inside that directory)
    import data                   # /root/reference/percivaltts/data.py
    assert os.path.dirname(os.path.abspath(data.__file__)) == REF and os.path.dirname(os.path.abspath(pt.__file__)) == REF
    return data, pt


def main():
    sys.path.insert(0, HERE)
    import data_scenarios
    data, pt = import_reference()
    with tempfile.TemporaryDirectory() as d:
        res = data_scenarios.run(data, pt, d)
    out = os.path.join(HERE, 'data_golden.npz')
    np.savez_compressed(out, **res)
    print('%s: %d entries, %d bytes' % (out, len(res), os.path.getsize(out)))


if __name__ == '__main__':
    main()
