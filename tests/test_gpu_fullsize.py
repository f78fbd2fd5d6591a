"""BASELINE.json's full-size configurations (C4: 1,000,000 x 2048, A = 32; C5: 20,000 x 4096, A = 64) checked through
size-independent properties — the oracle cannot run these shapes in seconds, the algebra can be verified exactly:

 * the weights have unit norm, R'P is unit upper triangular, the scores T = X R are mutually orthogonal and
   t_a't_a equals the tt the pass produced (p_a = X't_a / tt)  [Dayal-MacGregor invariants, SURVEY §4];
 * KERNEL_TYPE1 (streaming passes) == KERNEL_TYPE2 (DMMA X'X + SYMV): two different kernel paths, same model;
 * a fit is bit-reproducible run to run (no atomics anywhere);
 * LOO by Gram downdate == explicit refit on the N-1 remaining rows (C5 shape, a few folds);
 * rows 0..4095 of the device generator are bit-identical to the oracle's generator."""
import numpy as np
import pytest

from oracle import port
from pls_b200 import _capi
from tests.util import RTOL, assert_state_close, relerr

pytestmark = [pytest.mark.gpu, pytest.mark.slow]


@pytest.fixture(scope="module")
def ctx():
    c = _capi.Context(0)
    yield c
    c.close()


def test_c4_full_size_properties(ctx):
    seed, N, K, M, A = 20260004, 1_000_000, 2048, 8, 32
    ds = _capi.Dataset(ctx, synth=(seed, 0, N, K, M, 16))
    a = ds.fit(A, 0)
    W, P, R, Q, T = (a[k] for k in "WPRQT")
    assert np.all(np.isfinite(T)) and np.all(np.isfinite(R))
    assert np.allclose(np.linalg.norm(W, axis=0), 1.0, atol=1e-12)
    RtP = R.T @ P
    assert np.abs(np.tril(RtP, -1)).max() <= 1e-9 and np.abs(np.diag(RtP) - 1).max() <= 1e-9
    TtT = T.T @ T
    off = TtT - np.diag(np.diag(TtT))
    assert np.abs(off).max() <= 1e-9 * np.abs(np.diag(TtT)).max()
    # scores reproduce through the prediction path on a slice of the data: T[rows] == X[rows] R
    Xh, Yh = port.synth(seed, 4096, K, M)
    assert relerr(_capi.scores(ctx, R, A, Xh), T[:4096]) <= 1e-10
    # q_a = Y't_a / t_a't_a on the slice-independent identity: fitted + residual == Y on the slice
    E = _capi.fitted(ctx, R, Q, A, Xh, Yh); F = _capi.fitted(ctx, R, Q, A, Xh)
    assert relerr(E + F, Yh) <= 1e-12
    b = ds.fit(A, 1, want_T=False)                      # DMMA X'X path
    assert_state_close(a, b, keys="WPRQ", rtol=1e-8)
    Ba = port.coefficients(a["R"], a["Q"], A); Bb = port.coefficients(b["R"], b["Q"], A)
    assert relerr(Bb, Ba) <= RTOL
    c = ds.fit(A, 0)
    assert all(np.array_equal(a[k], c[k]) for k in "WPRQT")
    ds.close()


def test_c4_generator_head_is_the_oracles(ctx):
    ds = _capi.Dataset(ctx, synth=(20260004, 999_000, 1000, 2048, 8, 16))
    X, Y = ds.download()
    Xo, Yo = port.synth(20260004, 1_000_000, 2048, 8, row0=999_000, rows=1000)
    assert np.array_equal(X, Xo) and np.array_equal(Y, Yo)
    ds.close()


def test_c5_loo_downdate_equals_explicit_refit(ctx):
    seed, N, K, M, A = 20260005, 20_000, 4096, 4, 64
    ds = _capi.Dataset(ctx, synth=(seed, 0, N, K, M, 16))
    folds = np.array([[0], [7], [12345], [19999]], dtype=np.int64)
    E = ds.cv_folds(A, 0, folds)                        # [M][folds][A]
    X, Y = ds.download()
    ds.close()
    for f, (i,) in enumerate(folds):
        keep = np.ones(N, bool); keep[i] = False
        st = _capi.fit_host(ctx, X[keep], Y[keep], A, 0, want_T=False)        # streaming-pass fit on N-1 rows
        res = _capi.predict_residuals(ctx, st["R"], st["Q"], X[i:i + 1], Y[i:i + 1])   # [M][1][A]
        assert relerr(E[:, f, :], res[:, 0, :]) <= 1e-8, f
