"""
pctts_colsum_static: the bias-gradient column sum for hosts that cannot allocate device memory.

Some users of libpctts.so are handed their tensors and nothing else: a plugin living inside another framework's memory pool,
a caller whose device budget was fixed at start-up, a language binding with no CUDA allocator.  They cannot supply the scratch
workspace `pctts_colsum` takes, so `pctts_colsum_static` keeps that scratch in a fixed array inside the library (C <= 1024).
Here it is called through the C ABI and compared with `pctts_colsum` on the same inputs -- twice in a row with different
data and shapes, because every call reuses the one array and must not be disturbed by what the previous call left there.

The file sorts last on purpose (pytest -x stops at the first failure; a failure here must not hide the other tests), and the
test passes alone in a fresh process as well as after the rest of the suite.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_colsum_static_matches_colsum_on_two_consecutive_calls(cuda_ops):
    import torch
    from percivaltts_b200 import _lib
    lib = cuda_ops.lib
    stream = cuda_ops._stream()
    rng = np.random.RandomState(7)
    # (rows, C, accumulate): a wide and a narrow matrix, C = 4 takes the vectorised first stage; the second call is smaller
    # than the first, so stale partials of the first call sit right behind the slots it uses
    cases = [(4000, 256, 0), (333, 4, 1), (12000, 1024, 0), (57, 85, 1)]
    for rows, C, accumulate in cases:
        x = torch.from_numpy(rng.randn(rows, C).astype(np.float32)).cuda()
        start = torch.from_numpy(rng.randn(C).astype(np.float32)).cuda()
        want, got = start.clone(), start.clone()
        ws = cuda_ops.reduce_ws(rows, C)
        _lib.check(lib.pctts_colsum(x.data_ptr(), rows, C, want.data_ptr(), accumulate, ws.ptr, stream), 'pctts_colsum')
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        _lib.check(lib.pctts_colsum_static(x.data_ptr(), rows, C, got.data_ptr(), accumulate, stream), 'pctts_colsum_static')
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        # same kernels, same launch geometry, same summation order: identical bits (the suite's colsum tolerance is 1e-5)
        assert torch.equal(got, want), (rows, C, float((got - want).abs().max()))
        ref = x.double().sum(0) + (start.double() if accumulate else 0.0)
        scale = float(ref.abs().max()) + 1.0
        assert float((got.double() - ref).abs().max()) <= 1e-5 * scale, (rows, C)


def test_colsum_static_rejects_what_does_not_fit(cuda_ops):
    import torch
    lib = cuda_ops.lib
    x = torch.zeros(8, 1025, device='cuda')
    out = torch.full((1025,), 3.0, device='cuda')
    launches = lib.pctts_launch_count()
    assert lib.pctts_colsum_static(x.data_ptr(), 8, 1025, out.data_ptr(), 0, cuda_ops._stream()) == -3   # PCTTS_ELIMIT
    assert lib.pctts_colsum_static(0, 8, 4, out.data_ptr(), 0, cuda_ops._stream()) == -1                 # PCTTS_EINVAL
    assert lib.pctts_launch_count() == launches
    torch.cuda.synchronize()
    assert float(out.min()) == 3.0 and float(out.max()) == 3.0
