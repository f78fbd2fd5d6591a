"""Parity of the batched cross-validation engine (plsb_cv_folds, Gram-downdate + DMMA GEMM) with the
reference's cv_LOO / cv_LSO residuals (golden vectors from the unmodified reference) and with the CPU
oracle's explicit per-fold refits.  optimal_num_components must match exactly."""
import numpy as np
import pytest

from oracle import port
from pls_b200 import _capi
from pls_b200.model import MSE, RESS, Model, optimal_num_components, validation
from tests.util import RTOL, load_golden, relerr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = _capi.Context(0)
    yield c
    c.close()


def loo_idx(N):
    return np.arange(N, dtype=np.int64).reshape(-1, 1)


def test_c1_toy_loo_lso(ctx):
    g = load_golden("c1_toy")
    ds = _capi.Dataset(ctx, g["X"], g["Y"])
    loo = ds.cv_folds(2, 0, loo_idx(10))
    assert relerr(loo, g["loo"]) <= RTOL
    lso = ds.cv_folds(2, 0, g["lso_test"].astype(np.int64))
    assert relerr(lso, g["lso"]) <= RTOL
    assert relerr(_capi.validation(loo, 0), g["loo_ress"]) <= RTOL
    assert relerr(_capi.validation(lso, 1), g["lso_mse"]) <= RTOL
    assert np.array_equal(_capi.optimal_num_components(loo), g["loo_opt"])
    assert np.array_equal(_capi.optimal_num_components(lso), g["lso_opt"])
    ds.close()


@pytest.mark.parametrize("method", [0, 1])
def test_c2_nir_octane_loo(ctx, method):
    g = load_golden("c2_nir_octane")
    ds = _capi.Dataset(ctx, g["X"], g["Y"])
    loo = ds.cv_folds(10, method, loo_idx(60))
    assert relerr(loo, g["loo" if method == 0 else "loo_t2"]) <= RTOL
    assert relerr(_capi.validation(loo, 1), g["loo_mse"]) <= RTOL          # RMSEP within 1e-9
    assert np.array_equal(_capi.optimal_num_components(loo), g["loo_opt"])  # exactly (= 3)
    te = g["lso_test"].astype(np.int64)
    lso = ds.cv_folds(10, method, te)
    assert relerr(lso[:, :20 * 18, :], g["lso_first"]) <= RTOL
    assert relerr(_capi.validation(lso, 0), g["lso_ress"]) <= RTOL
    assert np.array_equal(_capi.optimal_num_components(lso), g["lso_opt"])
    ds.close()


def test_c3_nir_pls2_lso_1000_folds(ctx):
    g = load_golden("c3_nir_pls2")
    ds = _capi.Dataset(ctx, g["X"], g["Y"])
    te = g["lso_test"].astype(np.int64)
    lso = ds.cv_folds(10, 0, te)
    assert relerr(lso[:, :50 * 18, :], g["lso_first"]) <= RTOL
    assert relerr(lso[:, -18:, :], g["lso_last"]) <= RTOL
    assert relerr(_capi.validation(lso, 0), g["lso_ress"]) <= RTOL
    assert relerr(_capi.validation(lso, 1), g["lso_mse"]) <= RTOL
    assert np.array_equal(_capi.optimal_num_components(lso), g["lso_opt"])
    ds.close()


def test_c5s_loo(ctx):
    g = load_golden("c5s_loo")
    seed, N, K, M, A = (int(v) for v in g["meta"])
    ds = _capi.Dataset(ctx, synth=(seed, 0, N, K, M, 16))
    loo = ds.cv_folds(A, 0, loo_idx(N))
    assert relerr(loo, g["loo"]) <= RTOL
    assert relerr(_capi.validation(loo, 0), g["loo_ress"]) <= RTOL
    assert np.array_equal(_capi.optimal_num_components(loo), g["loo_opt"])
    ds.close()


@pytest.mark.parametrize("N,K,M,A,ts,nf", [(45, 11, 2, 5, 9, 30), (200, 77, 1, 6, 1, 200), (120, 300, 3, 5, 40, 7),
                                           (64, 130, 8, 4, 3, 50), (700, 20, 2, 4, 300, 3),
                                           # K >= 1024: the fused per-component fold kernel, every M variant
                                           (90, 1100, 3, 5, 1, 40), (60, 2049, 1, 4, 7, 12), (50, 4096, 8, 4, 2, 9),
                                           (70, 1024, 2, 6, 20, 6), (40, 1500, 5, 9, 1, 40),
                                           # 8 columns per thread (K <= 2048) with every M variant, ragged K
                                           (64, 1536, 8, 5, 2, 20), (48, 2048, 6, 4, 1, 30), (52, 1026, 4, 7, 5, 11),
                                           # outside the fused kernel's range: K > 4096, 8 < M <= 32 at K >= 1024
                                           (30, 4500, 2, 3, 1, 10), (44, 1100, 12, 4, 2, 9),
                                           # M > 32: per-fold scratch in global memory (step_wide.cuh)
                                           (80, 45, 36, 4, 3, 12), (60, 1100, 34, 3, 1, 10)])
def test_cv_matches_oracle_and_fold_sharding(ctx, N, K, M, A, ts, nf):
    X, Y = port.synth(500 + N + K, N, K, M, L=max(A, 4))
    rng = np.random.default_rng(N)
    te = np.stack([rng.permutation(N)[:ts] for _ in range(nf)]).astype(np.int64) if ts > 1 else loo_idx(N)[:nf]
    want = port.cv_folds(X, Y, A, 0, te)
    ds = _capi.Dataset(ctx, X, Y)
    got = ds.cv_folds(A, 0, te)
    assert relerr(got, want) <= RTOL
    # fold sharding: two calls covering disjoint fold ranges fill the same residual block
    out = np.zeros((M, A, nf * ts))
    h = nf // 2
    ds.cv_folds(A, 0, te, fold0=0, nfolds_local=h, out=out)
    ds.cv_folds(A, 0, te, fold0=h, nfolds_local=nf - h, out=out)
    assert np.array_equal(out.transpose(0, 2, 1), got)
    # small waves give the same answer as one big wave
    import os
    os.environ["PLSB_FOLD_BATCH"] = "3"
    try:
        assert np.array_equal(ds.cv_folds(A, 0, te), got)
    finally:
        del os.environ["PLSB_FOLD_BATCH"]
    # both regimes — one launch with the data set in shared memory (cv_small.cu, when it fits) and Gram downdates +
    # batched DMMA GEMM (cv.cu) — against the same explicit refits
    os.environ["PLSB_CV_NO_SMALL"] = "1"
    try:
        assert relerr(ds.cv_folds(A, 0, te), want) <= RTOL
    finally:
        del os.environ["PLSB_CV_NO_SMALL"]
    # PRESS reduced on the device (validation's column sums, pls.cpp:248-252), residual block kept there
    press = ds.cv_press(A, 0, te)
    assert relerr(press, (got ** 2).sum(axis=1)) <= 1e-13
    assert relerr(press, _capi.validation(got, 0)) <= 1e-13
    p0 = ds.cv_press(A, 0, te, fold0=0, nfolds_local=h); p1 = ds.cv_press(A, 0, te, fold0=h, nfolds_local=nf - h)
    assert relerr(p0 + p1, press) <= 1e-13
    ds.close()


@pytest.mark.parametrize("N,K,M,A", [(2600, 2200, 2, 4), (900, 37, 3, 5), (640, 1030, 1, 3)])
def test_two_lane_engine_equals_one_lane_and_the_oracle(ctx, N, K, M, A, monkeypatch):
    """>= 512 folds run as two waves in flight (the persistent, pre-tiled DMMA product of one lane under the TMA-ring-fed
    fold kernel of the other: cv_gemm.cu, cv.cu); one lane on the context's stream with the general GEMM is the A/B
    reference.  Folds are independent, every kernel is deterministic and neither product splits the reduction axis, so
    the two schedules must agree bit for bit; a sample of folds is checked against explicit refits by the oracle."""
    import os
    monkeypatch.setenv("PLSB_CV_NO_SMALL", "1")     # (the small shape would otherwise run as one launch out of shared memory)
    X, Y = port.synth(991, N, K, M, L=4)
    ds = _capi.Dataset(ctx, X, Y)
    te = loo_idx(N)
    two = ds.cv_folds(A, 0, te)
    os.environ["PLSB_CV_LANES"] = "1"
    try:
        one = ds.cv_folds(A, 0, te)
    finally:
        del os.environ["PLSB_CV_LANES"]
    assert np.array_equal(one, two)
    os.environ["PLSB_FOLD_BATCH"] = "300"       # several rounds of two waves, a ragged last one
    try:
        assert np.array_equal(ds.cv_folds(A, 0, te), two)
    finally:
        del os.environ["PLSB_FOLD_BATCH"]
    pick = np.array([0, 1, N // 2 - 1, N // 2, N // 2 + 107, N // 2 + 108, N - 1], dtype=np.int64)
    want = port.cv_folds(X, Y, A, 0, te[pick])
    assert relerr(two[:, pick, :], want) <= RTOL
    press = ds.cv_press(A, 0, te)
    assert relerr(press, (two ** 2).sum(axis=1)) <= 1e-13
    ds.close()


def test_two_lane_engine_with_several_held_out_rows(ctx, monkeypatch):
    """leave-some-out folds (3 held-out rows) through the two-lane engine: the downdate subtracts three rows per fold"""
    monkeypatch.setenv("PLSB_CV_NO_SMALL", "1")
    N, K, M, A, ts, nf = 500, 1300, 3, 4, 3, 600
    X, Y = port.synth(313, N, K, M, L=4)
    rng = np.random.default_rng(5)
    te = np.stack([rng.permutation(N)[:ts] for _ in range(nf)]).astype(np.int64)
    ds = _capi.Dataset(ctx, X, Y)
    got = ds.cv_folds(A, 0, te)                                   # [M][nf*ts][A]
    pick = np.array([0, 1, 299, 300, 599])
    want = port.cv_folds(X, Y, A, 0, te[pick])
    rows = (pick[:, None] * ts + np.arange(ts)[None, :]).ravel()
    assert relerr(got[:, rows, :], want) <= RTOL
    ds.close()


def test_cross_products_kept_on_the_data_set(ctx, monkeypatch):
    """plsb_dataset_build_gram: X'X and X'Y are built once and kept; cv calls (and fits) start from them."""
    monkeypatch.setenv("PLSB_CV_NO_SMALL", "1")
    N, K, M, A = 300, 140, 2, 5
    X, Y = port.synth(77, N, K, M, L=4)
    ds = _capi.Dataset(ctx, synth=(77, 0, N, K, M, 4))
    te = loo_idx(N)
    before = ds.cv_folds(A, 0, te)
    ds.build_gram()
    ctx.profile(True); ctx.profile_reset()
    after = ds.cv_folds(A, 0, te)
    assert ctx.profile_get("xtx")[1] == 0                 # not rebuilt by the call
    ctx.profile(False)
    assert np.array_equal(before, after)
    assert relerr(after[:, :40, :], port.cv_folds(X, Y, A, 0, te[:40])) <= RTOL
    got = ds.fit(A, 0)                                    # a fit from the kept X'X (KERNEL_TYPE2 recurrence) + T = X R
    want = port.fit(X, Y, A, 0)
    assert relerr(port.coefficients(got["R"], got["Q"], A), port.coefficients(want["R"], want["Q"], A)) <= RTOL
    ds.close()


def test_model_mirror_reads_like_the_reference(ctx):
    g = load_golden("c2_nir_octane")
    m = Model(g["X"], g["Y"], 0, 10, ctx=ctx)
    assert relerr(m.coefficients(), g["coef_A"]) <= RTOL
    assert relerr(m.fitted_values(g["X"]), g["fitted_A"]) <= RTOL
    assert relerr(np.array([m.SSE(g["X"], g["Y"], c) for c in range(1, 11)]), g["sse"]) <= RTOL
    assert np.abs(np.array([m.explained_variance(g["X"], g["Y"], c) for c in range(1, 11)]) - g["ev"]).max() <= 1e-9
    loo = m.cv_LOO()
    assert loo.method() == "LOO" and len(loo.errors()) == 1 and loo.errors()[0].shape == (60, 10)
    assert relerr(validation(loo, MSE), g["loo_mse"]) <= RTOL
    assert list(optimal_num_components(loo)) == [3]
    lso = m.cv_LSO(0.3, 600)
    assert relerr(validation(lso, RESS), g["lso_ress"]) <= RTOL
    assert np.array_equal(optimal_num_components(lso), g["lso_opt"])
    nd = m.cv_NEW_DATA(g["X"][:7], g["Y"][:7])
    assert nd.method() == "NEW DATA" and relerr(nd.array, g["new_data"]) <= RTOL
    assert m.loadingsX(3).shape == (401, 3) and m.loadingsY().shape == (1, 10)
