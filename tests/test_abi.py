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


def test_host_planning_calls_need_no_gpu():
    """The workspace / partition planners of the C ABI are host code: launch geometry of the demo shapes, checked without a device."""
    lib = _lib.load()
    # Conv1D k=21 425->256 forward, M = 4000: 66 tiles of 128 x 128 -> two K halves per tile, 2 x [4000, 256] floats of workspace
    assert lib.pctts_conv1d_tc_fwd_workspace(10, 400, 21, 425, 256) == 2 * 4000 * 256
    # its weight gradient: 84 x 2 = 168 tiles on 148 SMs -> 74 whole m-tiles, the other 10 (output rows 7906..8924) cut into 7 K ranges
    assert lib.pctts_conv1d_tc_wgrad_workspace(10, 400, 21, 425, 256) == 7 * (21 * 425 - (18 * 425 + 256)) * 256
    # critic post layer on the batched rows: 94 x 2 = 188 tiles, 13 k-blocks -> 3 ranges over rows 9472..11999
    assert lib.pctts_gemm_tc_tail_workspace(12000, 256, 772) == 3 * (12000 - 74 * 128) * 256
    assert lib.pctts_gemm_tc_tail_workspace(4000, 256, 256) == 0          # fewer tiles than SMs
    assert lib.pctts_gemm_tc_tail_workspace(12000, 256, 256) == 0         # 4 k-blocks: nothing to cut
    assert lib.pctts_gemm_tc_splits(256, 256, 4000) > 1 and lib.pctts_gemm_tc_splits(4000, 256, 256) == 1
    # BatchNorm statistics: partial slots of the column reduction and of the Conv2D epilogue
    assert lib.pctts_colreduce_nblk(10 * 400 * 129, 4) == 296 and lib.pctts_colreduce_nblk(4000, 256) == 296
    assert lib.pctts_colreduce_nblk(0, 4) == 0
    assert lib.pctts_conv2d_tc_stat_partials(10, 400, 129) == 148
    assert 0 < lib.pctts_conv2d_tc_stat_partials(2, 40, 33) < 148
    assert lib.pctts_conv2d_tc_supported(129, 4, 4, 5, 5, 2, 2) == 1 and lib.pctts_conv2d_tc_supported(129, 1, 4, 5, 5, 2, 2) == 0
    assert lib.pctts_peer_allreduce_bytes(8, 1024) > 2 * 1024 * 8 and lib.pctts_peer_allreduce_bytes(17, 1024) == -1
    # round 2: the chain kernel's cluster size per sample (demo: 8 CTAs of 50 frames; two operand planes fit 80 frames per CTA; a
    # sequence too long for 16 CTAs of shared memory is refused), the one-launch stack weight gradient's partial-sum workspace
    # (CTAs divided between the layers), the control block of the fused all-reduce + Adam, the sync-BN buffers incl. the push areas
    assert lib.pctts_conv2d_chain_supported(400, 129, 6) == 8 and lib.pctts_conv2d_chain_supported(50, 129, 6) == 1
    assert lib.pctts_conv2d_chain_supported(54, 129, 6) == 1 and lib.pctts_conv2d_chain_supported(55, 129, 6) == 2
    assert lib.pctts_conv2d_chain_supported(400, 129, 3) == 8 and lib.pctts_conv2d_chain_supported(4000, 129, 6) == 0
    assert lib.pctts_conv2d_chain_supported(0, 129, 6) == 0 and lib.pctts_conv2d_chain_supported(400, 129, 0) == 0
    # no cluster CTA may be left without frames: wide maps that would need 16 CTAs with an empty last one are refused
    assert lib.pctts_conv2d_chain_supported(225, 227, 6) == 0 and lib.pctts_conv2d_chain_supported(209, 243, 6) == 0
    assert lib.pctts_conv2d_chain_supported(224, 227, 6) == 8 and lib.pctts_conv2d_chain_supported(226, 227, 6) == 16
    for nterms in (6, 3, 1):
        for Fq in range(1, 520, 3):
            for T in range(1, 800, 7):
                nc = lib.pctts_conv2d_chain_supported(T, Fq, nterms)
                assert nc == 0 or (nc - 1) * ((T + nc - 1) // nc) < T, (T, Fq, nterms, nc)
    assert lib.pctts_conv2d_wgrad_multi_workspace(8) == 8 * 18 * 404 and lib.pctts_conv2d_wgrad_multi_workspace(1) == 148 * 404
    assert lib.pctts_conv2d_wgrad_multi_workspace(0) == 0 and lib.pctts_conv2d_wgrad_multi_workspace(10) == 0
    assert lib.pctts_grad_sync_ctl_bytes() == 512
    assert lib.pctts_bn_sync_bytes(8) > 2 * 32 * 8 * 128 * 8 and lib.pctts_bn_sync_bytes(17) == -1

