"""The C-ABI shared library builds for sm_100a, loads, and exports exactly what include/pctts.h declares
(no compute calls here: this container has no GPU)."""
import ctypes
import os
import subprocess

import pytest

from percivaltts_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_builds_loads_and_exports_every_declared_symbol():
    path = build.build()
    assert os.path.exists(path)
    protos = _lib.parse_header()
    assert len(protos) >= 30 and 'pctts_gemm_f32' in protos and 'pctts_pml_linbnd2fwbnd' in protos
    lib = ctypes.CDLL(path)
    for name in protos:
        assert hasattr(lib, name), name
    out = subprocess.run(['nm', '-D', '--defined-only', path], capture_output=True, text=True).stdout
    exported = set(ln.split()[-1] for ln in out.splitlines() if ' T pctts_' in ln)
    assert exported == set(protos), exported ^ set(protos)
    assert _lib.load().pctts_abi_version() == 1


def test_sass_is_sm100a_only():
    out = subprocess.run(['cuobjdump', '-lelf', build.LIBPATH], capture_output=True, text=True).stdout
    archs = set(ln.split('.')[-2] for ln in out.splitlines() if '.cubin' in ln)
    assert archs == {'sm_100a'}, archs


def test_product_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from percivaltts_b200.ops import CudaOps
    with pytest.raises(_lib.PcttsError):
        CudaOps()
    from percivaltts_b200 import sigproc
    from percivaltts_b200 import backend_b200
    backend_b200.set_ops(None)
    import numpy as np
    with pytest.raises(_lib.PcttsError):
        sigproc.linbnd2fwbnd(np.zeros((2, 2049), dtype=np.float32), 16000, 4096, 129)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'percivaltts_b200')
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith('.py') or fn.endswith('.cu') or fn.endswith('.cuh'):
                txt = open(os.path.join(dirpath, fn)).read()
                assert 'import oracle' not in txt and 'from oracle' not in txt and 'cpu_ops_double' not in txt.replace(
                    'tests/cpu_ops_double.py', ''), fn
