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


def test_extended_precision_fit_is_pinned_by_the_reference_goldens():
    """oracle/pls_oracle_ext.c (the long-double yardstick used where two correct FP64 fits differ by more than 1e-9) must
    agree with the reference's own results where conditioning is benign: the down-scaled synthetic configurations."""
    from tests.util import state_err
    for name in ("c4s", "c4s_odd", "c4s_m1"):
        g = load_golden(name)
        seed, N, K, M, A = (int(v) for v in g["meta"])
        X, Y = port.synth(seed, N, K, M)
        ext = port.fit_ext(X, Y, A)
        e, by = state_err(ext, {k: g.get("t1_" + k) for k in "WPRQT"})
        assert e <= 1e-10, (name, by)
        assert relerr(port.coefficients(ext["R"], ext["Q"], A), g["t1_coef_A"]) <= 1e-11


def test_full_size_fixtures_are_consistent():
    """tests/golden/c4_full.npz / c5_full.npz (the reference at BASELINE.json's full sizes): shapes, the recorded
    distances between independent evaluations, and coefficients == R Q' inside the fixture itself."""
    import json
    g = load_golden("c4_full")
    seed, N, K, M, A = (int(v) for v in g["meta"])
    assert (N, K, M, A) == (1_000_000, 2048, 8, 32) and g["W"].shape == (K, A) and g["T_sub"].shape == (192, A)
    assert relerr(port.coefficients(g["R"], g["Q"], A), g["coef_A"]) <= 1e-13
    assert relerr(port.coefficients(g["R"], g["Q"], 8), g["coef_8"]) <= 1e-13
    d = json.loads(str(g["distances_json"]))
    # the facts the GPU tests lean on: coefficients agree to ~3e-11 between every pair of evaluations, loadings only to
    # 0.5-1.2e-9, and the reference's own result is as far from the exact fit as any of the others
    for pair in ("oracle_type1_vs_reference_type1", "reference_type2_vs_reference_type1", "reference_type1_vs_extended"):
        assert d[pair]["coefficients"] <= 1e-10 and 1e-10 <= d[pair]["P"] <= 2e-9, (pair, d[pair])
    Xh, _ = port.synth(seed, 4, K, M)
    assert np.array_equal(Xh[:, :8], g["x_head"])
    g5 = load_golden("c5_full")
    seed, N, K, M, A = (int(v) for v in g5["meta"])
    assert (N, K, M, A) == (20_000, 4096, 4, 64) and g5["loo"].shape == (M, 8, A)
    d5 = json.loads(str(g5["distances_json"]))
    assert d5["oracle_loo_vs_reference"] <= 1e-9 and d5["reference_type1_vs_extended"]["P"] <= 1e-9
