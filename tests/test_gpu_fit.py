"""Parity of the CUDA fit / predict path (through the C-ABI) with the reference: committed golden
vectors produced by the unmodified reference, and the CPU oracle on seeded inputs.  Tolerances are the
north star's: 1e-9 relative (scores / loadings after sign normalisation)."""
import numpy as np
import pytest

from oracle import port
from pls_b200 import _capi
from tests.util import RTOL, assert_state_close, assert_state_close_or_exact, load_golden, relerr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = _capi.Context(0)
    yield c
    c.close()


def test_synth_is_bit_identical_to_oracle(ctx):
    for (seed, N, K, M, L, row0) in [(123, 1000, 77, 3, 16, 0), (20260004, 300, 256, 8, 16, 4096), (5, 33, 15, 1, 4, 7)]:
        ds = _capi.Dataset(ctx, synth=(seed, row0, N, K, M, L))
        X, Y = ds.download()
        Xo, Yo = port.synth(seed, row0 + N, K, M, L=L, row0=row0, rows=N)
        assert np.array_equal(X, Xo) and np.array_equal(Y, Yo)
        ds.close()


def test_upload_download_roundtrip(ctx):
    rng = np.random.default_rng(0)
    for N, K, M in [(1, 1, 1), (37, 15, 2), (513, 203, 3), (64, 1025, 8)]:
        X = np.asfortranarray(rng.standard_normal((N, K))); Y = np.asfortranarray(rng.standard_normal((N, M)))
        ds = _capi.Dataset(ctx, X, Y)
        X2, Y2 = ds.download()
        assert np.array_equal(X, X2) and np.array_equal(Y, Y2)
        ds.close()


@pytest.mark.parametrize("name,A", [("c1_toy", 2), ("c2_nir_octane", 10), ("c3_nir_pls2", 10)])
@pytest.mark.parametrize("method", [0, 1])
def test_fit_matches_reference_golden(ctx, name, A, method):
    g = load_golden(name)
    pre = "" if method == 0 else "t2_"
    ds = _capi.Dataset(ctx, g["X"], g["Y"])
    st = ds.fit(A, method)
    assert_state_close(st, {k: g.get(pre + k) for k in "WPRQT"})
    R, Q = st["R"], st["Q"]
    B = _capi.fitted(ctx, R, Q, A, np.eye(R.shape[0]))
    assert relerr(B, g[pre + "coef_A"]) <= RTOL
    assert relerr(_capi.fitted(ctx, R, Q, 1, np.eye(R.shape[0])), g[pre + "coef_1"]) <= RTOL
    assert relerr(_capi.fitted(ctx, R, Q, A, g["X"]), g[pre + "fitted_A"]) <= RTOL
    sse = np.array([(_capi.fitted(ctx, R, Q, c, g["X"], g["Y"]) ** 2).sum(axis=0) for c in range(1, A + 1)])
    assert relerr(sse, g[pre + "sse"]) <= RTOL
    sc = _capi.scores(ctx, R, A, g["X"])
    sgn = np.sign((sc * g[pre + "scores_A"]).sum(axis=0))
    assert relerr(sc * sgn, g[pre + "scores_A"]) <= RTOL
    ds.close()


@pytest.mark.parametrize("name", ["c4s", "c4s_odd", "c4s_m1"])
@pytest.mark.parametrize("method", [0, 1])
def test_fit_synthetic_matches_reference_golden(ctx, name, method):
    g = load_golden(name)
    seed, N, K, M, A = (int(v) for v in g["meta"])
    ds = _capi.Dataset(ctx, synth=(seed, 0, N, K, M, 16))
    pre = f"t{method + 1}_"
    st = ds.fit(A, method)
    assert_state_close(st, {k: g.get(pre + k) for k in "WPRQT"})
    X, _ = ds.download()
    assert np.array_equal(X[:4, :8], g["x_head"])
    assert relerr(_capi.fitted(ctx, st["R"], st["Q"], A, np.eye(K)), g[pre + "coef_A"]) <= RTOL
    assert relerr(_capi.fitted(ctx, st["R"], st["Q"], A, X[:64]), g[pre + "fitted_head"]) <= RTOL
    ds.close()


# every column-range of the pass-kernel dispatch, ragged N, M = 1 and M > 1
SHAPES = [(3, 2, 1, 1), (40, 7, 1, 5), (33, 12, 2, 6), (257, 64, 3, 6), (200, 100, 5, 9), (64, 130, 8, 5),
          (300, 255, 2, 7), (129, 401, 3, 8), (90, 513, 2, 6), (150, 1000, 4, 6), (70, 1500, 2, 5),
          (64, 2048, 8, 6), (50, 2500, 3, 5), (40, 4096, 4, 6), (2000, 31, 9, 9),
          (30, 5000, 2, 4), (25, 9001, 3, 4), (600, 4100, 12, 5),   # wide: the two-kernel fallback pass, M > 8
          (120, 50, 40, 5), (300, 600, 33, 4), (90, 70, 65, 3)]     # M > 32: the general step (step_wide.cuh)


@pytest.mark.parametrize("N,K,M,A", SHAPES)
@pytest.mark.parametrize("method", [0, 1])
def test_fit_matches_oracle(ctx, N, K, M, A, method):
    X, Y = port.synth(1000 + N + K, N, K, M, L=max(A, 4))
    want = port.fit(X, Y, A, method)
    got = _capi.fit_host(ctx, X, Y, A, method)
    assert_state_close_or_exact(got, want, X, Y, A)       # 1e-9 for every shape (W, P, R, Q, T per column)
    Bw = port.coefficients(want["R"], want["Q"], A)
    Bg = port.coefficients(got["R"], got["Q"], A)
    assert relerr(Bg, Bw) <= RTOL


def test_cross_products(ctx):
    X, Y = port.synth(99, 3000, 300, 5)
    ds = _capi.Dataset(ctx, X, Y)
    assert relerr(ds.xty(), X.T @ Y) <= 1e-12
    assert relerr(ds.xtx(), X.T @ X) <= 1e-12
    ds.close()


@pytest.mark.parametrize("N,K,M", [(1003, 700, 11), (5, 256, 1), (4099, 2047, 8), (777, 1024, 16), (3, 1999, 2)])
def test_xty_dmma_path(ctx, N, K, M):
    """X'Y (pls.cpp:396) through the DMMA + TMA kernel (xty.cu): ragged row tails, M not a multiple of 8."""
    X, Y = port.synth(4000 + N, N, K, M)
    ds = _capi.Dataset(ctx, X, Y)
    S = ds.xty()
    assert relerr(S, X.T @ Y) <= 1e-12
    assert np.array_equal(S, ds.xty())          # reproducible bit for bit
    ds.close()


@pytest.mark.parametrize("N,K,M,A", [(3000, 300, 3, 8), (1500, 77, 1, 6), (2048, 512, 8, 10), (5000, 1030, 2, 5)])
def test_fit_from_gram_accumulated_during_upload(ctx, N, K, M, A, monkeypatch):
    """Host data uploaded in row blocks with X'X accumulated on the tensor cores under the transfer (dataset.cu):
    the fit that follows runs from X'X and forms T = X R with the DMMA scores kernel.  Same model as plsr
    (pls.cpp:390-437, KERNEL_TYPE1 request), ragged last block, odd K."""
    monkeypatch.setenv("PLSB_UPLOAD_GRAM", "1")
    X, Y = port.synth(7000 + N + K, N, K, M, L=max(A, 4))
    want = port.fit(X, Y, A, 0)
    ds = _capi.Dataset(ctx, X, Y)
    ctx.profile(True); ctx.profile_reset()
    got = ds.fit(A, 0)
    assert ctx.profile_get("pass_fused")[1] == 0 and ctx.profile_get("scores")[1] >= 1     # no streaming pass ran
    ctx.profile(False)
    assert_state_close_or_exact(got, want, X, Y, A)
    assert relerr(port.coefficients(got["R"], got["Q"], A), port.coefficients(want["R"], want["Q"], A)) <= RTOL
    assert relerr(ds.xtx(), X.T @ X) <= 1e-12
    Xd, Yd = ds.download()
    assert np.array_equal(Xd, X) and np.array_equal(Yd, Y)
    monkeypatch.setenv("PLSB_FIT_NO_GRAM", "1")                     # the streaming-pass path on the same data set
    again = ds.fit(A, 0)
    assert_state_close_or_exact(again, want, X, Y, A)
    ds.zscore()                                                     # X changes in place: the Gram matrix is dropped
    monkeypatch.delenv("PLSB_FIT_NO_GRAM")
    z = ds.fit(A, 0)
    Xz = (X - X.mean(0)) / X.std(0, ddof=1); Yz = (Y - Y.mean(0)) / Y.std(0, ddof=1)
    wz = port.fit(np.asfortranarray(Xz), np.asfortranarray(Yz), A, 0)
    assert relerr(port.coefficients(z["R"], z["Q"], A), port.coefficients(wz["R"], wz["Q"], A)) <= RTOL
    ds.close()


@pytest.mark.parametrize("N,K", [(20011, 2048), (9000, 1500), (700, 300)])
def test_pass_occupancy_does_not_change_the_bits(ctx, N, K, monkeypatch):
    """The fused pass deals its tiles to a fixed number of virtual CTAs, so running one CTA per SM (deep ring) or two
    (shallow ring) — chosen by the size of the data (pass.cu: launch_fused_pass) — gives bit-identical models."""
    ds = _capi.Dataset(ctx, synth=(99, 0, N, K, 3, 8))
    monkeypatch.setenv("PLSB_PASS_CTAS_PER_SM", "2")
    a = ds.fit(6, 0)
    monkeypatch.setenv("PLSB_PASS_CTAS_PER_SM", "1")
    b = ds.fit(6, 0)
    monkeypatch.setenv("PLSB_PASS_STAGES", "2")
    c = ds.fit(6, 0)
    assert all(np.array_equal(a[k], b[k]) and np.array_equal(a[k], c[k]) for k in "WPRQT")
    if K > 1024:   # the large-stream configuration (one 512-thread CTA per SM, two row groups), forced on a small shape
        monkeypatch.delenv("PLSB_PASS_CTAS_PER_SM"); monkeypatch.delenv("PLSB_PASS_STAGES")
        monkeypatch.setenv("PLSB_PASS_RB", "12")
        d = ds.fit(6, 0)
        assert_state_close(d, a, rtol=1e-11)
        X, Y = ds.download()
        assert_state_close_or_exact(d, port.fit(X, Y, 6, 0), X, Y, 6)
    ds.close()


def test_type1_equals_type2_and_invariants(ctx):
    ds = _capi.Dataset(ctx, synth=(7, 0, 5000, 384, 4, 16))
    a = ds.fit(12, 0); b = ds.fit(12, 1)
    assert_state_close(a, b, keys="WPRQ")
    R, P, W, T = a["R"], a["P"], a["W"], a["T"]
    assert np.allclose(np.linalg.norm(W, axis=0), 1.0, atol=1e-12)
    RtP = R.T @ P                                    # unit upper triangular
    assert np.allclose(np.tril(RtP, -1), 0, atol=1e-9) and np.allclose(np.diag(RtP), 1, atol=1e-9)
    TtT = T.T @ T
    off = TtT - np.diag(np.diag(TtT))
    assert np.abs(off).max() <= 1e-9 * np.abs(np.diag(TtT)).max()
    # run-to-run reproducibility (no atomics)
    c = ds.fit(12, 0)
    assert all(np.array_equal(a[k], c[k]) for k in "WPRQT")
    ds.close()


@pytest.mark.parametrize("Nn,K,A", [(1, 7, 1), (130, 2049, 33), (257, 512, 8), (1000, 33, 17), (129, 4100, 40)])
def test_scores_dmma_shapes(ctx, Nn, K, A):
    """S = X_new R (pls.cpp:439-442) through the DMMA kernel (scores.cu): ragged row tiles, odd K (zero-padded
    leading dimension), component counts that are not multiples of 8 and exceed one 32-component launch."""
    rng = np.random.default_rng(Nn + K + A)
    Xn = np.asfortranarray(rng.standard_normal((Nn, K)))
    R = np.asfortranarray(rng.standard_normal((K, A)))
    for comp in sorted({1, A // 2 or 1, A}):
        assert relerr(_capi.scores(ctx, R, comp, Xn), Xn @ R[:, :comp]) <= 1e-13


def test_predict_helpers_match_oracle(ctx):
    X, Y = port.synth(31, 500, 90, 3)
    st = port.fit(X, Y, 7, 0)
    Xn, Yn = port.synth(32, 123, 90, 3)
    R, Q = st["R"], st["Q"]
    assert relerr(_capi.scores(ctx, R, 5, Xn), port.scores(R, 5, Xn)) <= RTOL
    assert relerr(_capi.fitted(ctx, R, Q, 4, Xn), port.fitted_values(R, Q, 4, Xn)) <= RTOL
    assert relerr(_capi.fitted(ctx, R, Q, 7, Xn, Yn), port.residuals(R, Q, 7, Xn, Yn)) <= RTOL
    assert relerr(_capi.predict_residuals(ctx, R, Q, Xn, Yn), port.predict_residuals(R, Q, Xn, Yn)) <= RTOL
    g = load_golden("c2_nir_octane")
    ds = _capi.Dataset(ctx, g["X"], g["Y"])
    s2 = ds.fit(10, 0)
    nd = _capi.predict_residuals(ctx, s2["R"], s2["Q"], g["X"][:7], g["Y"][:7])
    assert relerr(nd, g["new_data"]) <= RTOL
    ds.close()


def test_argument_errors(ctx):
    X, Y = port.synth(1, 20, 5, 1)
    ds = _capi.Dataset(ctx, X, Y)
    with pytest.raises(_capi.PlsbError):
        ds.fit(0, 0)
    with pytest.raises(_capi.PlsbError):
        ds.fit(2, 7)
    ds.close()


def test_column_statistics_and_z_scores(ctx):
    """SST / colwise_stdev / colwise_z_scores on the device (pls.cpp:69-111) against the reference's golden values
    and numpy's two-pass formulas; the zero-variance NaN of the reference is reproduced."""
    for name in ("c1_toy", "c2_nir_octane"):
        g = load_golden(name)
        mean, sst, sd = _capi.colstats(ctx, g["X_raw"])
        if "x_sst" in g:
            assert relerr(sst, g["x_sst"]) <= 1e-12 and relerr(sd, g["x_stdev"]) <= 1e-12
        assert relerr(_capi.colwise_z_scores(ctx, g["X_raw"]), g["X"]) <= 1e-12
        assert relerr(_capi.colwise_z_scores(ctx, g["Y_raw"]), g["Y"]) <= 1e-12
        assert relerr(_capi.colwise_z_scores(ctx, g["X_raw"], mean, sd), g["X"]) <= 1e-12
    rng = np.random.default_rng(5)
    for N, K in [(2, 1), (37, 15), (1001, 203), (64, 1025)]:
        X = rng.standard_normal((N, K)) * rng.uniform(0.1, 10, K) + rng.uniform(-5, 5, K)
        mean, sst, sd = _capi.colstats(ctx, X)
        mu = X.mean(axis=0)
        assert relerr(mean, mu) <= 1e-13
        assert relerr(sst, ((X - mu) ** 2).sum(axis=0)) <= 1e-12
        assert relerr(sd, X.std(axis=0, ddof=1)) <= 1e-12
        assert relerr(_capi.colstats(ctx, X, means=np.zeros(K))[1], (X ** 2).sum(axis=0)) <= 1e-12   # SST(mat, means)
        ds = _capi.Dataset(ctx, X, X[:, :1])
        xm, xs, ym, ys = ds.zscore()
        Z, _ = ds.download()
        assert relerr(Z, (X - mu) / X.std(axis=0, ddof=1)) <= 1e-12 and relerr(xs, sd) <= 1e-12
        ds.close()
    assert np.array_equal(_capi.colstats(ctx, np.ones((1, 3)))[1], np.zeros(3))       # N < 2: zeros (pls.cpp:70-71)
    Xc = rng.standard_normal((10, 4)); Xc[:, 2] = 3.0
    Z = _capi.colwise_z_scores(ctx, Xc)
    assert np.isnan(Z[:, 2]).all() and np.isfinite(Z[:, [0, 1, 3]]).all()


def test_complex_scores_download_small_and_staged_and_after_trim(ctx):
    """plsb_model_download_scores(as_complex=1) fills the drop-in's complex-typed Model::T (pls.h:251-254).  Below 1 MB it
    widens in place in the caller's buffer (no pinned staging, no host threads); above, it goes through the staging
    pipeline — and plsb_ctx_trim, which frees that pipeline's buffers, must leave the next large transfer working."""
    rng = np.random.default_rng(31)
    for N, A in [(50, 3), (1, 1), (70001, 2), (50, 3)]:        # 1.2 KB, 8 B, 1.12 MB (staged), small again
        X = rng.standard_normal((N, 12)); Y = rng.standard_normal((N, 2))
        ds = _capi.Dataset(ctx, X, Y)
        out = ds.fit_resident(A, 0, download="TC")
        assert np.array_equal(out["C"].real, out["T"]) and not out["C"].imag.any()
        assert np.abs(out["T"]).max() > 0
        ds.close()
        ctx.trim()


@pytest.mark.parametrize("method", [0, 1])
@pytest.mark.parametrize("N,K,M", [(8, 12, 2), (5, 9, 1), (6, 20, 3)])
def test_components_past_the_numerical_rank(ctx, N, K, M, method):
    """The 3-argument constructor asks for A = K components (pls.cpp:356-359); with K > N the data have rank N and the
    reference's components N+1..K are finite garbage (a deflated cross product of rounding noise, normalised: loadings
    up to 1e15).  The first N components are still the model, held to 1e-9; the rest must stay finite like the reference's."""
    X, Y = port.synth(77 + N, N, K, M, L=3)
    want = port.fit(X, Y, K, method)
    got = _capi.fit_host(ctx, X, Y, K, method)
    head = lambda st: {k: (None if st.get(k) is None else st[k][:, :N]) for k in "WPRQT"}
    assert_state_close(head(got), head(want))
    for k in "WPRQ":
        assert np.isfinite(got[k]).all(), k
