"""Multi-GPU paths; skipped on a single-GPU box.

 * row-sharded single fit: every GPU holds a row shard, K*M / K+1 partial sums are all-reduced (NCCL);
   results must equal the single-GPU fit to 1e-12 and the oracle to 1e-9.
 * fold-sharded cross-validation: folds split over GPUs, no communication; identical residual blocks.
 * the same two through the C++ drop-in layer (PLSB_DEVICES / PLSB_SHARD_MIN_ROWS)."""
import os
import threading

import numpy as np
import pytest

from oracle import port
from pls_b200 import _capi
from tests.util import RTOL, assert_state_close, load_golden, relerr

pytestmark = pytest.mark.gpu


def ngpu():
    try:
        return _capi.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(ngpu() < 2, reason="needs >= 2 GPUs")


def shard_bounds(N, G):
    return [i * N // G for i in range(G + 1)]


@needs2
@pytest.mark.parametrize("method", [0, 1])
def test_row_sharded_fit_matches_single_gpu(method):
    G = min(ngpu(), 4)
    seed, N, K, M, A = 777, 30011, 300, 5, 12
    X, Y = port.synth(seed, N, K, M)
    ctxs = [_capi.Context(i) for i in range(G)]
    uid = _capi.Context.unique_id()
    b = shard_bounds(N, G)
    out = [None] * G

    def rank(i):
        ctxs[i].comm_init(i, G, uid)
        ds = _capi.Dataset(ctxs[i], X[b[i]:b[i + 1]], Y[b[i]:b[i + 1]], sharded=True)
        out[i] = ds.fit(A, method)
        ds.close()

    th = [threading.Thread(target=rank, args=(i,)) for i in range(G)]
    [t.start() for t in th]; [t.join() for t in th]
    assert all(o is not None for o in out)
    single = _capi.Dataset(ctxs[0], X, Y).fit(A, method)
    for i in range(G):                                    # K-sized results are replicated on every rank
        for k in "WPRQ":
            assert relerr(out[i][k], single[k]) <= 1e-11, (i, k)
    if method == 0:
        T = np.vstack([o["T"] for o in out])
        assert relerr(T, single["T"]) <= 1e-11
    want = port.fit(X, Y, A, method)
    got = dict(out[0]); got["T"] = np.vstack([o["T"] for o in out]) if method == 0 else None
    assert_state_close(got, want)
    Bg = port.coefficients(got["R"], got["Q"], A); Bw = port.coefficients(want["R"], want["Q"], A)
    assert relerr(Bg, Bw) <= RTOL
    [c.close() for c in ctxs]


@needs2
@pytest.mark.parametrize("mixed", [False, True])
def test_row_sharded_fit_from_gram_accumulated_during_upload(mixed, monkeypatch):
    """Row shards uploaded with X'X / X'Y accumulated under the transfer (dataset.cu): the partial Gram matrices are
    all-reduced once and the fit runs from the sum.  mixed: one rank has no Gram matrix (device-generated shard) —
    the ranks must agree on the streaming-pass path instead of dead-locking in different collectives."""
    G = 2
    seed, N, K, M, A = 4242, 6000, 300, 3, 9
    X, Y = port.synth(seed, N, K, M)
    monkeypatch.setenv("PLSB_UPLOAD_GRAM", "1")
    ctxs = [_capi.Context(i) for i in range(G)]
    uid = _capi.Context.unique_id()
    b = shard_bounds(N, G)
    out = [None] * G

    def rank(i):
        ctxs[i].comm_init(i, G, uid)
        if mixed and i == 1:
            ds = _capi.Dataset(ctxs[i], synth=(seed, b[i], b[i + 1] - b[i], K, M, 16), sharded=True)
        else:
            ds = _capi.Dataset(ctxs[i], X[b[i]:b[i + 1]], Y[b[i]:b[i + 1]], sharded=True)
        ctxs[i].profile(True); ctxs[i].profile_reset()
        out[i] = ds.fit(A, 0)
        out[i]["passes"] = ctxs[i].profile_get("pass_fused")[1]
        out[i] = dict(out[i], again=ds.fit(A, 0))          # second fit: the Gram sum must not be reduced twice
        ds.close()

    th = [threading.Thread(target=rank, args=(i,)) for i in range(G)]
    [t.start() for t in th]; [t.join() for t in th]
    assert all(o is not None for o in out)
    assert all((o["passes"] > 0) == mixed for o in out)
    want = port.fit(X, Y, A, 0)
    for key in ("first", "again"):
        got = {k: (out[0] if key == "first" else out[0]["again"])[k] for k in "WPRQ"}
        got["T"] = np.vstack([(o if key == "first" else o["again"])["T"] for o in out])
        assert_state_close(got, want)
        assert relerr(port.coefficients(got["R"], got["Q"], A), port.coefficients(want["R"], want["Q"], A)) <= RTOL
    [c.close() for c in ctxs]


@needs2
def test_fold_sharded_cv_matches_single_gpu():
    G = min(ngpu(), 4)
    g = load_golden("c3_nir_pls2")
    te = g["lso_test"].astype(np.int64)
    ctxs = [_capi.Context(i) for i in range(G)]
    E = np.zeros((3, 10, te.size))
    fb = shard_bounds(te.shape[0], G)

    def rank(i):
        ds = _capi.Dataset(ctxs[i], g["X"], g["Y"])
        ds.cv_folds(10, 0, te, fold0=fb[i], nfolds_local=fb[i + 1] - fb[i], out=E)
        ds.close()

    th = [threading.Thread(target=rank, args=(i,)) for i in range(G)]
    [t.start() for t in th]; [t.join() for t in th]
    Em = E.transpose(0, 2, 1)
    single = _capi.Dataset(ctxs[0], g["X"], g["Y"]).cv_folds(10, 0, te)
    assert np.array_equal(Em, single)                     # no communication, same kernels: bit-identical
    assert relerr(Em[:, :g["lso_first"].shape[1]], g["lso_first"]) <= RTOL
    assert np.array_equal(_capi.optimal_num_components(Em), g["lso_opt"])
    [c.close() for c in ctxs]


@needs2
def test_cross_products_built_across_ranks_for_fold_sharded_cv(monkeypatch):
    """plsb_dataset_build_gram(across_ranks): every rank holds the whole data set (folds are sharded) but reads only its
    row slice for X'X / X'Y; the partial matrices are all-reduced.  Residuals agree with the single-GPU call (summation
    order differs: 1e-11) and with the oracle's explicit refits."""
    monkeypatch.setenv("PLSB_CV_NO_SMALL", "1")
    G = 2
    seed, N, K, M, A = 31, 1200, 260, 2, 6
    X, Y = port.synth(seed, N, K, M, L=4)
    te = np.arange(N, dtype=np.int64).reshape(-1, 1)
    ctxs = [_capi.Context(i) for i in range(G)]
    uid = _capi.Context.unique_id()
    E = np.zeros((M, A, N))
    fb = shard_bounds(N, G)
    ok = [False] * G

    def rank(i):
        ctxs[i].comm_init(i, G, uid)
        ds = _capi.Dataset(ctxs[i], synth=(seed, 0, N, K, M, 4))
        ds.build_gram(across_ranks=True)
        ctxs[i].profile(True); ctxs[i].profile_reset()
        ds.cv_folds(A, 0, te, fold0=fb[i], nfolds_local=fb[i + 1] - fb[i], out=E)
        ok[i] = ctxs[i].profile_get("xtx")[1] == 0
        ctxs[i].profile(False)
        ds.close()

    th = [threading.Thread(target=rank, args=(i,)) for i in range(G)]
    [t.start() for t in th]; [t.join() for t in th]
    assert all(ok)
    Em = E.transpose(0, 2, 1)
    single = _capi.Dataset(ctxs[0], synth=(seed, 0, N, K, M, 4)).cv_folds(A, 0, te)
    assert relerr(Em, single) <= 1e-11
    pick = np.array([0, 599, 600, 1199])
    assert relerr(Em[:, pick, :], port.cv_folds(X, Y, A, 0, te[pick])) <= RTOL
    [c.close() for c in ctxs]


@needs2
def test_cpp_layer_multi_gpu(tmp_path):
    from tests.test_gpu_cpp_api import API_CHECK, run_api_check, stacked
    if not os.path.exists(API_CHECK):
        pytest.skip("api_check not built")
    g = load_golden("c3_nir_pls2")
    devs = ",".join(str(i) for i in range(min(ngpu(), 4)))
    one = run_api_check(tmp_path / "a", g["X"], g["Y"], 10, 0, 0.3, 200) if (tmp_path / "a").mkdir() is None else None
    many = run_api_check(tmp_path / "b", g["X"], g["Y"], 10, 0, 0.3, 200,
                         env={"PLSB_DEVICES": devs, "PLSB_SHARD_MIN_ROWS": "1"}) if (tmp_path / "b").mkdir() is None else None
    for k in ("coef_A", "fitted_A", "sse"):
        assert relerr(many[k], one[k]) <= 1e-11, k
        assert relerr(many[k], g[k]) <= RTOL, k
    # predictions are row-sharded over the GPUs too (scores, residuals, explained variance, cv_NEW_DATA)
    for k in ("scores_A", "resid_A", "ev", "R"):
        assert relerr(many[k], one[k]) <= 1e-11, k
    assert relerr(stacked(many, "newdata", 3), stacked(one, "newdata", 3)) <= 1e-11
    assert np.array_equal(stacked(many, "lso", 3), stacked(one, "lso", 3))
    assert np.array_equal(stacked(many, "loo", 3), stacked(one, "loo", 3))
