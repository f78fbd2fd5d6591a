"""Pin the CPU oracle (oracle/pls_oracle.c) against golden vectors produced by the UNMODIFIED
reference (tests/golden/make_golden.py).  Runs everywhere; no GPU, no /root/reference needed."""
import numpy as np
import pytest

from oracle import port
from tests.util import RTOL, assert_state_close, load_golden, relerr


@pytest.mark.parametrize("name,A", [("c1_toy", 2), ("c2_nir_octane", 10), ("c3_nir_pls2", 10)])
@pytest.mark.parametrize("method", [0, 1])
def test_fit_matches_reference(name, A, method):
    g = load_golden(name)
    pre = "" if method == 0 else "t2_"
    st = port.fit(g["X"], g["Y"], A, method)
    want = {k: g.get(pre + k) for k in "WPRQT"}
    assert_state_close(st, want)
    assert relerr(port.coefficients(st["R"], st["Q"], A), g[pre + "coef_A"]) <= RTOL
    assert relerr(port.coefficients(st["R"], st["Q"], 1), g[pre + "coef_1"]) <= RTOL
    assert relerr(port.fitted_values(st["R"], st["Q"], A, g["X"]), g[pre + "fitted_A"]) <= RTOL
    sse = np.array([port.sse(st["R"], st["Q"], c, g["X"], g["Y"]) for c in range(1, A + 1)])
    assert relerr(sse, g[pre + "sse"]) <= 1e-8     # SSE at c=A is tiny vs c=1: matrix-relative
    ev = np.array([port.explained_variance(st["R"], st["Q"], c, g["X"], g["Y"]) for c in range(1, A + 1)])
    assert np.abs(ev - g[pre + "ev"]).max() <= 1e-9
    sc = port.scores(st["R"], A, g["X"])
    sgn = np.sign((sc * g[pre + "scores_A"]).sum(axis=0))
    assert relerr(sc * sgn, g[pre + "scores_A"]) <= RTOL


@pytest.mark.parametrize("name", ["c4s", "c4s_odd", "c4s_m1"])
@pytest.mark.parametrize("method", [0, 1])
def test_fit_synthetic_matches_reference(name, method):
    g = load_golden(name)
    seed, N, K, M, A = (int(v) for v in g["meta"])
    X, Y = port.synth(seed, N, K, M)
    assert np.array_equal(X[:4, :8], g["x_head"])            # generator unchanged since the goldens
    pre = f"t{method + 1}_"
    st = port.fit(X, Y, A, method)
    assert_state_close(st, {k: g.get(pre + k) for k in "WPRQT"})
    assert relerr(port.coefficients(st["R"], st["Q"], A), g[pre + "coef_A"]) <= RTOL
    assert relerr(port.fitted_values(st["R"], st["Q"], A, X[:64]), g[pre + "fitted_head"]) <= RTOL


def test_zscores_and_stats():
    g = load_golden("c1_toy")
    assert relerr(port.colwise_z_scores(g["X_raw"]), g["X"]) <= 1e-12
    assert relerr(port.colwise_stdev(g["X_raw"]), g["x_stdev"]) <= 1e-13
    assert relerr(port.sst(g["X_raw"]), g["x_sst"]) <= 1e-13
    s = load_golden("stats")
    assert np.allclose([port.normalcdf(z) for z in s["z"]], s["cdf"], rtol=0, atol=1e-15)
    assert np.allclose([port.wilcoxon(a, b) for a, b in zip(s["e1"], s["e2"])], s["wil"], rtol=0, atol=1e-14)


def test_c1_loo_lso():
    g = load_golden("c1_toy")
    te = np.arange(10).reshape(-1, 1)
    loo = port.cv_folds(g["X"], g["Y"], 2, 0, te)
    assert relerr(loo, g["loo"]) <= RTOL
    lso = port.cv_folds(g["X"], g["Y"], 2, 0, g["lso_test"].astype(np.int64), g["lso_train"].astype(np.int64))
    assert relerr(lso, g["lso"]) <= RTOL
    assert relerr(port.validation(loo, 0), g["loo_ress"]) <= RTOL
    assert relerr(port.validation(lso, 1), g["lso_mse"]) <= RTOL
    assert np.array_equal(port.optimal_num_components(loo), g["loo_opt"])
    assert np.array_equal(port.optimal_num_components(lso), g["lso_opt"])
    # known-answer text values of SURVEY Appendix C (captured from the reference CLI)
    assert np.allclose(np.sqrt(g["loo_mse"]), [[0.791811, 0.419116], [1.03315, 1.05394]], rtol=2e-6)
    assert list(g["loo_opt"]) == [2, 1]


def test_c2_loo_optimal_components():
    g = load_golden("c2_nir_octane")
    te = np.arange(60).reshape(-1, 1)
    for method, key in ((0, "loo"), (1, "loo_t2")):
        loo = port.cv_folds(g["X"], g["Y"], 10, method, te)
        assert relerr(loo, g[key]) <= RTOL
    loo = port.cv_folds(g["X"], g["Y"], 10, 0, te)
    assert np.array_equal(port.optimal_num_components(loo), g["loo_opt"])
    assert list(g["loo_opt"]) == [3]
    assert relerr(port.validation(loo, 1), g["loo_mse"]) <= RTOL
    p = [port.wilcoxon(loo[0][:, 6], loo[0][:, c]) for c in range(6)]
    assert np.allclose(p, g["wilcoxon_7_vs"], rtol=0, atol=1e-12)
    # cv_NEW_DATA layout (pls.cpp:494-510)
    st = port.fit(g["X"], g["Y"], 10, 0)
    nd = port.predict_residuals(st["R"], st["Q"], g["X"][:7], g["Y"][:7])
    assert relerr(nd, g["new_data"]) <= RTOL


def test_c3_lso_pls2():
    g = load_golden("c3_nir_pls2")
    tr = g["lso_train"].astype(np.int64); te = g["lso_test"].astype(np.int64)
    lso = port.cv_folds(g["X"], g["Y"], 10, 0, te, tr)
    assert relerr(lso[:, :50 * 18, :], g["lso_first"]) <= RTOL
    assert relerr(lso[:, -18:, :], g["lso_last"]) <= RTOL
    assert relerr(port.validation(lso, 0), g["lso_ress"]) <= RTOL
    assert np.array_equal(port.optimal_num_components(lso), g["lso_opt"])


def test_c5s_loo():
    g = load_golden("c5s_loo")
    seed, N, K, M, A = (int(v) for v in g["meta"])
    X, Y = port.synth(seed, N, K, M)
    te = np.arange(N).reshape(-1, 1)
    loo = port.cv_folds(X, Y, A, 0, te)
    assert relerr(loo, g["loo"]) <= RTOL
    assert np.array_equal(port.optimal_num_components(loo), g["loo_opt"])


def test_synth_properties():
    X, Y = port.synth(123, 2000, 32, 3)
    Xb, Yb = port.synth(123, 2000, 32, 3, row0=500, rows=100)
    assert np.array_equal(X[500:600], Xb) and np.array_equal(Y[500:600], Yb)   # row blocks are consistent
    u = np.array([port.lib().orc_synth_u(9, i) for i in range(20000)])
    assert abs(u.mean()) < 0.02 and abs(u.var() - 0.25) < 0.01
