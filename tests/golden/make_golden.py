#!/usr/bin/env python
"""Generate the golden vectors in tests/golden/ from the UNMODIFIED reference.

Runs only where /root/reference is mounted (the build container): it drives
oracle/_ref/libplsref.so (the reference compiled by oracle/Makefile) and oracle/_ref/PLS_ref (the
reference CLI) and stores inputs + outputs as small .npz / .txt fixtures so that the parity tests
can run on the GPU box, where the reference does not exist.

    python tests/golden/make_golden.py        # ~1-2 min (the reference's own cv_LOO on nir is 33 s)
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import port, ref  # noqa: E402

REF = "/root/reference"


def save(name, **kw):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **kw)
    print(f"{name}.npz  {os.path.getsize(path)/1024:.1f} KB")


def model_dump(X, Y, A, method, comps=None):
    m = ref.Model(X, Y, A, method)
    assert m.max_imag() == 0.0
    st = m.state()
    out = {f"{k}": v for k, v in st.items() if v is not None}
    comps = comps or list(range(1, A + 1))
    out["sse"] = np.array([m.sse(X, Y, c) for c in comps])
    out["ev"] = np.array([m.explained_variance(X, Y, c) for c in comps])
    out["coef_A"] = m.coefficients(A)
    out["coef_1"] = m.coefficients(1)
    out["fitted_A"] = m.fitted_values(X, A)
    out["scores_A"] = m.scores(X, A)
    return m, out


def main():
    assert ref.available(), "needs /root/reference"
    toyX = ref.read_matrix_file(f"{REF}/toyX.csv"); toyY = ref.read_matrix_file(f"{REF}/toyY.csv")
    nir = ref.read_matrix_file(f"{REF}/nir.csv"); octane = ref.read_matrix_file(f"{REF}/octane.csv")

    # ---- C1: toy, CLI semantics (z-scored, type 1, A=2; main.cpp:21-41) ----------------------
    X = ref.colwise_z_scores(toyX); Y = ref.colwise_z_scores(toyY)
    m, d = model_dump(X, Y, 2, 0)
    loo = m.cv_loo()                                  # reference's own driver
    lso = m.cv_lso(0.3, 100)                          # mt19937 default seed 5489 (main.cpp:38)
    tr, te = ref.lso_folds(10, 0.3, 100)
    lso_fast = ref.cv_folds(X, Y, 2, 0, te, tr)
    assert np.array_equal(lso, lso_fast), "inner-A LSO must be bit-identical to the reference's inner-K"
    te_loo = np.arange(10, dtype=np.int64).reshape(-1, 1)
    assert np.array_equal(loo, ref.cv_folds(X, Y, 2, 0, te_loo)), "inner-A LOO must be bit-identical"
    m2, d2 = model_dump(X, Y, 2, 1)
    save("c1_toy", X_raw=toyX, Y_raw=toyY, X=X, Y=Y, **d, **{"t2_" + k: v for k, v in d2.items()},
         loo=loo, lso=lso, lso_train=tr.astype(np.int16), lso_test=te.astype(np.int16),
         loo_ress=ref.validation(loo, 0), loo_mse=ref.validation(loo, 1), loo_opt=ref.optimal_num_components(loo),
         lso_ress=ref.validation(lso, 0), lso_mse=ref.validation(lso, 1), lso_opt=ref.optimal_num_components(lso),
         x_stdev=ref.colwise_stdev(toyX), x_sst=ref.sst(toyX))
    r = subprocess.run([ref.CLI, f"{REF}/toyX.csv", f"{REF}/toyY.csv", "2"], capture_output=True, text=True)
    with open(os.path.join(HERE, "c1_cli_stderr.txt"), "w") as f:
        f.write(r.stderr)
    assert r.returncode == 0 and r.stdout == ""

    # ---- C2: nir/octane, A=10, LOO + optimal_num_components ----------------------------------
    X = ref.colwise_z_scores(nir); Y = ref.colwise_z_scores(octane)
    m, d = model_dump(X, Y, 10, 0)
    print("C2: reference cv_LOO as written (inner A = K = 401) ...", flush=True)
    loo = m.cv_loo()
    te_loo = np.arange(60, dtype=np.int64).reshape(-1, 1)
    assert np.array_equal(loo, ref.cv_folds(X, Y, 10, 0, te_loo))
    tr, te = ref.lso_folds(60, 0.3, 600)
    lso = ref.cv_folds(X, Y, 10, 0, te, tr)
    assert np.array_equal(m.cv_lso(0.3, 3), lso[:, :3 * 18, :])       # the real driver, first 3 folds
    m2, d2 = model_dump(X, Y, 10, 1)
    loo2 = ref.cv_folds(X, Y, 10, 1, te_loo)
    newd = m.cv_new_data(X[:7], Y[:7])
    save("c2_nir_octane", X_raw=nir, Y_raw=octane, X=X, Y=Y, **d, **{"t2_" + k: v for k, v in d2.items()},
         loo=loo, loo_t2=loo2, loo_ress=ref.validation(loo, 0), loo_mse=ref.validation(loo, 1),
         loo_opt=ref.optimal_num_components(loo),
         lso_train=tr.astype(np.int16), lso_test=te.astype(np.int16),
         lso_ress=ref.validation(lso, 0), lso_opt=ref.optimal_num_components(lso), lso_first=lso[:, :20 * 18, :],
         new_data=newd,
         wilcoxon_7_vs=np.array([ref.wilcoxon(loo[0][:, 6], loo[0][:, c]) for c in range(6)]))

    # ---- C3: nir PLS2: Y3 = [octane | nir[:,100] | nir[:,300]], 1000 LSO folds --
    Y3 = np.column_stack([octane[:, 0], nir[:, 100], nir[:, 300]])
    X3 = np.delete(nir, [100, 300], axis=1)
    X = ref.colwise_z_scores(X3); Y = ref.colwise_z_scores(Y3)
    m, d = model_dump(X, Y, 10, 0)
    tr, te = ref.lso_folds(60, 0.3, 1000)
    lso = ref.cv_folds(X, Y, 10, 0, te, tr)
    assert np.array_equal(m.cv_lso(0.3, 2), lso[:, :2 * 18, :])
    m2, d2 = model_dump(X, Y, 10, 1)
    save("c3_nir_pls2", X=X, Y=Y, **d, **{"t2_" + k: v for k, v in d2.items()},
         lso_train=tr.astype(np.int16), lso_test=te.astype(np.int16),
         lso_ress=ref.validation(lso, 0), lso_mse=ref.validation(lso, 1), lso_opt=ref.optimal_num_components(lso),
         lso_first=lso[:, :50 * 18, :], lso_last=lso[:, -18:, :])

    # ---- synthetic, C4-shaped, down-scaled (inputs regenerated from the seed at test time) -----
    for tag, (seed, N, K, M, A) in {"c4s": (20260004, 4096, 256, 8, 16), "c4s_odd": (20260104, 1531, 203, 3, 9),
                                     "c4s_m1": (20260204, 900, 77, 1, 12)}.items():
        X, Y = port.synth(seed, N, K, M)
        out = {}
        for meth in (0, 1):
            _, d = model_dump(X, Y, A, meth, comps=[1, A // 2, A])
            out.update({f"t{meth + 1}_{k}": v for k, v in d.items() if k not in ("fitted_A", "scores_A")})
            out[f"t{meth + 1}_fitted_head"] = d["fitted_A"][:64]
        save(tag, meta=np.array([seed, N, K, M, A]), x_head=X[:4, :8].copy(), **out)

    # ---- synthetic, C5-shaped, down-scaled: full LOO ---------------------------------------------
    seed, N, K, M, A = 20260005, 384, 96, 4, 12
    X, Y = port.synth(seed, N, K, M)
    te_loo = np.arange(N, dtype=np.int64).reshape(-1, 1)
    loo = ref.cv_folds(X, Y, A, 0, te_loo)
    loo2 = ref.cv_folds(X, Y, A, 1, te_loo)
    save("c5s_loo", meta=np.array([seed, N, K, M, A]), loo=loo, loo_t2=loo2, loo_ress=ref.validation(loo, 0),
         loo_opt=ref.optimal_num_components(loo))

    # ---- statistics known answers ---------------------------------------------------------------
    rng = np.random.default_rng(7)
    zs = np.linspace(-6, 6, 49)
    e1 = rng.standard_normal((8, 37)); e2 = rng.standard_normal((8, 37)) * 1.3
    save("stats", z=zs, cdf=np.array([ref.normalcdf(z) for z in zs]), e1=e1, e2=e2,
         wil=np.array([ref.wilcoxon(a, b) for a, b in zip(e1, e2)]))


if __name__ == "__main__":
    main()
