"""
percivaltts_b200/data.py and the configuration helpers against the REFERENCE's own code: tests/golden/data_golden.npz holds what
/root/reference/percivaltts/data.py and percivaltts.py (unmodified, run in the build container by
tests/golden/make_data_golden.py) returned on the seeded scenario of tests/golden/data_scenarios.py; the same scenario runs here
on this repository's modules.  Arrays -- batches, crops, masks, and the draw that follows each call on numpy's global RNG (the
number and order of draws is part of the reference's repeatability) -- must be identical bit for bit; the scalar costs within
1e-6 (the reference adds float32 partial sums).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, 'golden'))


def test_data_layer_matches_the_reference_outputs(tmp_path):
    import data_scenarios
    from percivaltts_b200 import data, percivaltts
    gold = np.load(os.path.join(HERE, 'golden', 'data_golden.npz'))
    state = np.random.get_state()
    try:
        got = data_scenarios.run(data, percivaltts, str(tmp_path))
    finally:
        np.random.set_state(state)
    assert set(got) == set(gold.files), set(got) ^ set(gold.files)
    assert len(gold.files) > 200
    for name in gold.files:
        want, have = gold[name], got[name]
        if name.startswith('cost.'):
            assert np.allclose(have, want, rtol=1e-6, atol=0), (name, have, want)
        else:
            assert have.shape == want.shape, (name, have.shape, want.shape)
            assert have.dtype == want.dtype, (name, have.dtype, want.dtype)
            assert np.array_equal(have, want), name
