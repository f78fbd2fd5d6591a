"""The C++ drop-in layer (include/PLS/pls.h + pls_b200/libpls.so) and the `PLS` command line tool on the
GPU, checked against golden vectors produced by the unmodified reference (tests/golden/make_golden.py).

tests/cpp/api_check is a small program written the way a user of the reference writes code
(PLS::Model, cv_LOO, validation, optimal_num_components ...); it dumps every result as a PLSBMAT1 file.
The binaries are built in the build container (Eigen headers are not in this repo) and travel with the
snapshot."""
import json
import os
import re
import struct
import subprocess

import numpy as np
import pytest

from tests.util import RTOL, col_relerr, load_golden, relerr

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
API_CHECK = os.path.join(ROOT, "tests", "cpp", "api_check")
CLI = os.path.join(ROOT, "pls_b200", "PLS")


def write_bin(path, a):
    a = np.asfortranarray(np.atleast_2d(a), dtype=np.float64)
    with open(path, "wb") as f:
        f.write(b"PLSBMAT1" + struct.pack("<QQ", a.shape[0], a.shape[1]))
        f.write(a.tobytes(order="F"))


def read_bin(path):
    with open(path, "rb") as f:
        assert f.read(8) == b"PLSBMAT1"
        r, c = struct.unpack("<QQ", f.read(16))
        return np.frombuffer(f.read(), dtype=np.float64).reshape((r, c), order="F")


def run_api_check(tmp_path, X, Y, A, method, frac, trials, zscore=False, env=None):
    write_bin(tmp_path / "X.bin", X); write_bin(tmp_path / "Y.bin", Y)
    out = tmp_path / "out"; out.mkdir(exist_ok=True)
    cmd = [API_CHECK, str(tmp_path / "X.bin"), str(tmp_path / "Y.bin"), str(A), str(method), str(out), str(frac), str(trials)]
    if zscore:
        cmd.append("zscore")
    e = dict(os.environ); e.update(env or {})
    r = subprocess.run(cmd, capture_output=True, text=True, env=e, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return {f[:-4]: read_bin(out / f) for f in os.listdir(out)}


def sign_fix(a, ref):
    """flip columns of `a` to the sign convention of `ref` (scores / loadings are defined up to sign)"""
    sg = np.sign((a * ref).sum(axis=0)); sg[sg == 0] = 1
    return a * sg[None, :]


def stacked(d, name, M):
    return np.stack([d[f"{name}_{m}"] for m in range(M)])


@pytest.mark.skipif(not os.path.exists(API_CHECK), reason="tests/cpp/api_check not built (needs Eigen headers)")
@pytest.mark.parametrize("name,A,method,trials", [("c1_toy", 2, 0, 100), ("c1_toy", 2, 1, 0), ("c2_nir_octane", 10, 0, 600),
                                                   ("c2_nir_octane", 10, 1, 0), ("c3_nir_pls2", 10, 0, 1000)])
def test_cpp_model_api_matches_reference(tmp_path, name, A, method, trials):
    g = load_golden(name)
    raw = "X_raw" in g
    d = run_api_check(tmp_path, g["X_raw"] if raw else g["X"], g["Y_raw"] if raw else g["Y"], A, method, 0.3, trials, zscore=raw)
    pre = "t2_" if method == 1 else ""
    M = g["Y"].shape[1]
    if raw:
        assert relerr(d["Xz"], g["X"]) <= 1e-12 and relerr(d["Yz"], g["Y"]) <= 1e-12
        if "x_stdev" in g:
            assert relerr(d["x_stdev"].ravel(), g["x_stdev"]) <= 1e-12 and relerr(d["x_sst"].ravel(), g["x_sst"]) <= 1e-12
    for k in ("P", "Q", "R", "scores_A"):
        assert col_relerr(sign_fix(d[k], g[pre + k]), g[pre + k]) <= RTOL, k
    for k in ("coef_A", "coef_1", "fitted_A"):
        assert relerr(d[k], g[pre + k]) <= RTOL, k
    assert relerr(d["resid_A"], g["Y"] - g[pre + "fitted_A"]) <= RTOL
    assert relerr(d["sse"], g[pre + "sse"]) <= RTOL and relerr(d["ev"], g[pre + "ev"]) <= RTOL
    assert relerr(d["alias_coef_A"], g[pre + "coef_A"]) <= RTOL
    loo = stacked(d, "loo", M)
    gl = g["loo_t2"] if (method == 1 and "loo_t2" in g) else g.get("loo")
    if gl is not None and (method == 0 or "loo_t2" in g):
        assert relerr(loo, gl) <= RTOL
    if method == 0 and "loo_ress" in g:
        assert relerr(d["loo_ress"], g["loo_ress"]) <= RTOL and relerr(d["loo_mse"], g["loo_mse"]) <= RTOL
        assert np.array_equal(d["loo_opt"].ravel().astype(int), g["loo_opt"].astype(int))   # exactly
        assert relerr(d["alias_loo_ress"], g["loo_ress"]) <= RTOL
    if trials:
        lso = stacked(d, "lso", M)
        if "lso" in g:
            assert relerr(lso, g["lso"]) <= RTOL
        if "lso_first" in g:
            n = g["lso_first"].shape[1]
            assert relerr(lso[:, :n], g["lso_first"]) <= RTOL
        if "lso_last" in g:
            n = g["lso_last"].shape[1]
            assert relerr(lso[:, -n:], g["lso_last"]) <= RTOL
        assert relerr(d["lso_ress"], g["lso_ress"]) <= RTOL
        assert np.array_equal(d["lso_opt"].ravel().astype(int), g["lso_opt"].astype(int))
        # cv_LSO takes the caller's generator by reference and must leave it where the reference does (pls.cpp:512, 534)
        want_rng = json.load(open(os.path.join(ROOT, "tests", "golden", "lso_rng_after.json")))[f"{name}:{trials}"]
        assert int(d["rng_after"].ravel()[0]) == want_rng
    if "new_data" in g and method == 0:
        assert relerr(stacked(d, "newdata", M), g["new_data"]) <= RTOL
    # refit on row-reversed data: the regression is invariant to the order of the observations
    assert relerr(d["refit_coef_A"], g[pre + "coef_A"]) <= RTOL


@pytest.mark.skipif(not os.path.exists(API_CHECK), reason="tests/cpp/api_check not built (needs Eigen headers)")
def test_cpp_model_survives_set_devices(tmp_path):
    """PLS::set_devices between the constructor's fit and cv_LOO / cv_LSO: the engine lets go of the contexts the model's
    device copy lives on (they are reference-counted by their data sets) and the copy is carried over to the new ones."""
    g = load_golden("c2_nir_octane")
    d = run_api_check(tmp_path, g["X_raw"], g["Y_raw"], 10, 0, 0.3, 600, zscore=True, env={"PLSB_TEST_RECONFIGURE": "1"})
    assert relerr(stacked(d, "loo", 1), g["loo"]) <= RTOL
    assert np.array_equal(d["loo_opt"].ravel().astype(int), g["loo_opt"].astype(int))
    assert relerr(d["lso_ress"], g["lso_ress"]) <= RTOL
    assert relerr(d["coef_A"], g["coef_A"]) <= RTOL


NUM = re.compile(r"[-+]?(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?")


@pytest.mark.skipif(not os.path.exists(CLI), reason="pls_b200/PLS not built (needs Eigen headers)")
def test_cli_toy_matches_reference_text(tmp_path):
    """config 1: `PLS toyX.csv toyY.csv 2` — same report on stderr as the reference CLI, to the 6 significant
    digits it prints (columns are sign-normalised: the reference's eigenvector sign is arbitrary)."""
    g = load_golden("c1_toy")
    np.savetxt(tmp_path / "toyX.csv", g["X_raw"], delimiter=",", fmt="%.17g")
    np.savetxt(tmp_path / "toyY.csv", g["Y_raw"], delimiter=",", fmt="%.17g")
    r = subprocess.run([CLI, str(tmp_path / "toyX.csv"), str(tmp_path / "toyY.csv"), "2"], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0 and r.stdout == "", r.stderr[-2000:]
    want = open(os.path.join(ROOT, "tests", "golden", "c1_cli_stderr.txt")).read()
    gl, wl = r.stderr.strip().splitlines(), want.strip().splitlines()
    assert len(gl) == len(wl)
    section, flips = None, {}
    for a, b in zip(gl, wl):
        if re.fullmatch(r"[A-Za-z]+:", b.strip()):
            assert a.strip() == b.strip()
            section, first = b.strip(), True
            continue
        assert NUM.sub("#", a).split() == NUM.sub("#", b).split(), (a, b)      # same words and layout of numbers
        na = np.array([float(x) for x in NUM.findall(a)]); nb = np.array([float(x) for x in NUM.findall(b)])
        if section in ("P:", "W:", "R:", "Q:", "T:") and "(" in b:
            na, nb = na[0::2], nb[0::2]          # real parts of "(re,im)"
            if section not in flips:             # per-component sign taken from the section's first row
                flips[section] = np.sign(na * nb)
            na = na * flips[section]
        if "Optimal number" in b or re.fullmatch(r"\d+", b.strip()):
            assert a.strip() == b.strip()
            continue
        assert np.allclose(na, nb, rtol=2e-5, atol=2e-6), (a, b)


def test_cli_usage_exit_code():
    if not os.path.exists(CLI):
        pytest.skip("pls_b200/PLS not built")
    r = subprocess.run([CLI], capture_output=True, text=True)
    assert r.returncode == 100 and "Usage" in r.stderr
