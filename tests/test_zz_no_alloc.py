"""plsb_colstats_static — the entry point for callers that cannot allocate device memory — against plsb_colstats.

Some users of this library live inside a program that owns the GPU's memory (an application-wide arena that leaves
the driver allocator nothing to give), or sit in a loop where a cudaMalloc would stall every stream on the device, or
must prove a fixed memory budget up front.  For them libplsb.so exports one column-statistics call that allocates
nothing: inputs and outputs are device buffers the caller already has, and the partial sums go to a fixed
4 MiB __device__ array inside the library (at most 8192 columns).  Here torch plays that caller: it owns every
buffer and the stream, and the library is only given raw addresses.

The file sorts last on purpose (`pytest -x` stops at the first failure, so nothing hides behind this test), and it
must pass both alone in a fresh process — where its first call does the per-device set-up — and after the rest
of the suite has already used the library in this process.
"""
import numpy as np
import pytest

from pls_b200 import _capi
from tests.util import relerr

pytestmark = pytest.mark.gpu

TOL = 1e-12   # what test_column_statistics_and_z_scores asks of plsb_colstats itself


def _static_stats(torch, Xd, N, K, out, stream):
    """One call through the C-ABI.  Xd: (K, ldx) contiguous torch tensor = column-major N x K with leading dimension ldx."""
    _capi.colstats_static(Xd.device.index, stream.cuda_stream, Xd.data_ptr(), Xd.shape[1], N, K,
                          out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr())
    stream.synchronize()
    return [o[:K].cpu().numpy() for o in out]


def test_static_scratch_colstats_matches_allocating_entry_point():
    import torch
    dev = torch.device("cuda", 0)
    ctx = _capi.Context(0)
    rng = np.random.default_rng(23)
    stream = torch.cuda.Stream(device=dev)
    # the caller's buffers, sized once for the widest case: reused by every call below, never handed to the library to grow
    out = [torch.full((8192,), float("nan"), dtype=torch.float64, device=dev) for _ in range(3)]
    try:
        # narrow after wide on purpose: the second call of each pair and the narrower shapes find the previous call's
        # partial sums still in the shared array and must not be affected by them
        for N, K, pad in [(1001, 203, 0), (5000, 2048, 7), (37, 15, 3), (2, 1, 0), (300, 8192, 0)]:
            X = rng.standard_normal((N, K)) * rng.uniform(0.1, 10, K) + rng.uniform(-5, 5, K)
            mean, sst, sd = _capi.colstats(ctx, X)                       # the existing, allocating entry point
            ldx = N + pad
            Xh = np.full((K, ldx), np.nan)                                # padding rows must never be read
            Xh[:, :N] = X.T
            Xd = torch.from_numpy(Xh).to(dev)
            torch.cuda.synchronize(dev)                                  # the upload ran on torch's default stream
            first = _static_stats(torch, Xd, N, K, out, stream)
            second = _static_stats(torch, Xd, N, K, out, stream)         # two calls in a row on the same array
            for got in (first, second):
                assert relerr(got[0], mean) <= TOL and relerr(got[1], sst) <= TOL and relerr(got[2], sd) <= TOL
            assert all(np.array_equal(a, b) for a, b in zip(first, second)), "fixed summation order: same bits every call"
        # outputs are optional; the means then stay inside the library's array
        only_sd = torch.zeros(8192, dtype=torch.float64, device=dev)
        _capi.colstats_static(0, stream.cuda_stream, Xd.data_ptr(), Xd.shape[1], N, K, None, None, only_sd.data_ptr())
        stream.synchronize()
        assert relerr(only_sd[:K].cpu().numpy(), sd) <= TOL
        # fewer than two rows: SST is zero, as in the reference (pls.cpp:70-71) and in plsb_colstats
        one = torch.ones((3, 1), dtype=torch.float64, device=dev)
        got = _static_stats(torch, one, 1, 3, out, stream)
        assert np.array_equal(got[0], np.ones(3)) and np.array_equal(got[1], np.zeros(3))
        assert np.array_equal(_capi.colstats(ctx, np.ones((1, 3)))[1], got[1])
        # beyond the array's capacity: refused with a message, nothing launched, outputs untouched
        before = out[0].clone()
        with pytest.raises(_capi.PlsbError, match="8192"):
            _capi.colstats_static(0, stream.cuda_stream, Xd.data_ptr(), 1, 1, 8193, out[0].data_ptr(), None, None)
        stream.synchronize()
        assert torch.equal(before, out[0])
    finally:
        ctx.close()
