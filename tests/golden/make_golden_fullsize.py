#!/usr/bin/env python
"""Full-size golden vectors (BASELINE.json configs 4 and 5) from the UNMODIFIED reference.

Runs only where /root/reference is mounted (the build container; ~50 GB of host memory at the peak, ~20 min):

  C4  synthetic 1,000,000 x 2048, Y x 8, 32 components: the reference's own Model(X, Y, KERNEL_TYPE1, 32)
      (src/pls.cpp:340-353 -> plsr 390-437), single thread as built.  Stored: coefficients(32) and a few prefixes,
      W, P, R, Q, the rows {0..63, 500000..500063, last 64} of T, fitted values and residuals of the first 64 rows.
      Beside them the DISTANCES between independent FP64 evaluations of the same model on the same bits — the
      reference's KERNEL_TYPE2 (its -fopenmp build, X'X through Eigen's threaded GEMM) and this repo's C oracle
      (another summation order) against the reference's KERNEL_TYPE1 — which is what "conditioning of the weights
      W" means in numbers (tests/test_gpu_fullsize.py states its W bound from these).
  C5  synthetic 20,000 x 4096, Y x 4, 64 components: the reference's full fit, and 8 leave-one-out folds refitted by
      the reference's plsr on the 19,999 remaining rows (src/pls.cpp:469-491 with inner A = 64; the first A components
      of its inner A = K fits are bit-identical, asserted in make_golden.py on the small configurations).

    python tests/golden/make_golden_fullsize.py [c4] [c5] [ext]     (ext: only add the extended-precision yardstick)
"""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import port, ref  # noqa: E402
from tests.util import col_relerr, relerr, sign_normalise  # noqa: E402

C4 = dict(seed=20260004, N=1_000_000, K=2048, M=8, A=32)
C5 = dict(seed=20260005, N=20_000, K=4096, M=4, A=64)
C4_TROWS = np.r_[0:64, 500_000:500_064, 999_936:1_000_000]
C5_FOLDS = np.array([0, 1, 7, 5000, 9999, 12345, 15000, 19999], dtype=np.int64)
C5_COLS = np.array([0, 1, 2, 3, 7, 15, 31, 47, 63])


def save(name, **kw):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **kw)
    print(f"{name}.npz  {os.path.getsize(path)/1024:.1f} KB", flush=True)


def distances(got, want, A):
    g = sign_normalise({k: got.get(k) for k in "WPRQT"}); w = sign_normalise({k: want.get(k) for k in "WPRQT"})
    d = {k: col_relerr(g[k], w[k]) for k in "WPRQT" if g.get(k) is not None and w.get(k) is not None}
    d["coefficients"] = relerr(port.coefficients(got["R"], got["Q"], A), port.coefficients(want["R"], want["Q"], A))
    return d


def c4():
    c = C4
    t0 = time.time()
    X, Y = port.synth(c["seed"], c["N"], c["K"], c["M"])
    print(f"C4 synth {time.time()-t0:.0f} s", flush=True)
    t0 = time.time()
    m = ref.Model(X, Y, c["A"], 0)
    t_ref = time.time() - t0
    print(f"C4 reference KERNEL_TYPE1 fit {t_ref:.0f} s", flush=True)
    assert m.max_imag() == 0.0
    st = m.state()
    head = X[:64].copy(order="F")
    out = dict(meta=np.array([c["seed"], c["N"], c["K"], c["M"], c["A"]]), W=st["W"], P=st["P"], R=st["R"], Q=st["Q"],
               T_rows=C4_TROWS, T_sub=st["T"][C4_TROWS], coef_A=m.coefficients(c["A"]),
               coef_1=m.coefficients(1), coef_8=m.coefficients(8), coef_16=m.coefficients(16),
               fitted_head=m.fitted_values(head, c["A"]), resid_head=m.residuals(head, Y[:64], c["A"]),
               scores_head=m.scores(head, c["A"]), x_head=X[:4, :8].copy())
    m.close(); del m
    dist = {"reference_type1_seconds_1_thread": t_ref}
    # the C oracle (tests' CPU restatement) at full size: pins the oracle at C4 and gives one distance
    t0 = time.time()
    o = port.fit(X, Y, c["A"], 0)
    dist["oracle_type1_seconds"] = time.time() - t0
    dist["oracle_type1_vs_reference_type1"] = distances(o, st, c["A"])
    print("C4 oracle:", dist["oracle_type1_vs_reference_type1"], flush=True)
    del o
    st.pop("T")
    if ref.omp_available():
        t0 = time.time()
        m2 = ref.Model(X, Y, c["A"], 1, omp=True)
        dist["reference_type2_openmp_seconds"] = time.time() - t0
        s2 = m2.state()
        out["t2_coef_A"] = m2.coefficients(c["A"])
        m2.close(); del m2
        dist["reference_type2_vs_reference_type1"] = distances(s2, st, c["A"])
        print("C4 reference type 2:", dist["reference_type2_vs_reference_type1"], flush=True)
    out["distances_json"] = np.array(json.dumps(dist))
    save("c4_full", **out)
    json.dump(dist, open(os.path.join(HERE, "c4_full_distances.json"), "w"), indent=1)


def c5():
    c = C5
    X, Y = port.synth(c["seed"], c["N"], c["K"], c["M"])
    t0 = time.time()
    m = ref.Model(X, Y, c["A"], 0)
    t_ref = time.time() - t0
    print(f"C5 reference KERNEL_TYPE1 fit {t_ref:.0f} s", flush=True)
    st = m.state()
    head = X[:64].copy(order="F")
    out = dict(meta=np.array([c["seed"], c["N"], c["K"], c["M"], c["A"]]), cols=C5_COLS, Q=st["Q"],
               W_cols=st["W"][:, C5_COLS], P_cols=st["P"][:, C5_COLS], R_cols=st["R"][:, C5_COLS],
               T_cols_head=st["T"][:256][:, C5_COLS], coef_A=m.coefficients(c["A"]), coef_32=m.coefficients(32),
               fitted_head=m.fitted_values(head, c["A"]), x_head=X[:4, :8].copy())
    # sign convention of the stored columns: the tests normalise by the largest |r| entry, which needs all of R's column
    R = st["R"]
    out["R_argmax"] = np.abs(R).argmax(axis=0)[C5_COLS]
    m.close(); del m
    dist = {"reference_type1_seconds_1_thread": t_ref}
    o = port.fit(X, Y, c["A"], 0)
    dist["oracle_type1_vs_reference_type1"] = distances(o, st, c["A"])
    print("C5 oracle:", dist["oracle_type1_vs_reference_type1"], flush=True)
    # leave-one-out folds by the reference's own plsr on the remaining rows, one fold per host thread
    t0 = time.time()
    with ThreadPoolExecutor(len(C5_FOLDS)) as ex:
        res = list(ex.map(lambda i: ref.cv_folds(X, Y, c["A"], 0, np.array([[i]], dtype=np.int64)), C5_FOLDS))
    dist["reference_loo_folds_seconds"] = time.time() - t0
    out["loo_folds"] = C5_FOLDS
    out["loo"] = np.concatenate(res, axis=1)          # [M][folds][A]
    lo = port.cv_folds(X, Y, c["A"], 0, C5_FOLDS[:2].reshape(-1, 1))
    dist["oracle_loo_vs_reference"] = relerr(lo, out["loo"][:, :2, :])
    print("C5 LOO folds done", dist["reference_loo_folds_seconds"], dist["oracle_loo_vs_reference"], flush=True)
    out["distances_json"] = np.array(json.dumps(dist))
    save("c5_full", **out)
    json.dump(dist, open(os.path.join(HERE, "c5_full_distances.json"), "w"), indent=1)


def ext(tag, c, trows):
    """Add the extended-precision fit (oracle/pls_oracle_ext.c) to an existing fixture: ext_* arrays + distances of the
    reference's FP64 result from it."""
    path = os.path.join(HERE, tag + ".npz")
    with np.load(path) as z:
        d = {k: z[k] for k in z.files}
    X, Y = port.synth(c["seed"], c["N"], c["K"], c["M"])
    t0 = time.time()
    e = port.fit_ext(X, Y, c["A"])
    dist = json.loads(str(d["distances_json"]))
    dist["ext_seconds"] = time.time() - t0
    A = c["A"]
    if "W" in d:      # C4 fixture: full W, P, R, Q and a row subset of T
        refst = dict(W=d["W"], P=d["P"], R=d["R"], Q=d["Q"], T=d["T_sub"])
        est = dict(W=e["W"], P=e["P"], R=e["R"], Q=e["Q"], T=e["T"][trows])
        dist["reference_type1_vs_extended"] = distances(refst, est, A)
        d.update(ext_W=e["W"], ext_P=e["P"], ext_R=e["R"], ext_Q=e["Q"], ext_T_sub=e["T"][trows],
                 ext_coef_A=port.coefficients(e["R"], e["Q"], A))
    else:             # C5 fixture: a column subset
        cols = d["cols"]
        sg_r = np.sign(d["R_cols"][d["R_argmax"], np.arange(len(cols))])
        sg_e = np.sign(e["R"][d["R_argmax"], cols])
        fl = (sg_r * sg_e)[None, :]
        dist["reference_type1_vs_extended"] = {
            "W": col_relerr(e["W"][:, cols] * fl, d["W_cols"]), "P": col_relerr(e["P"][:, cols] * fl, d["P_cols"]),
            "R": col_relerr(e["R"][:, cols] * fl, d["R_cols"]),
            "coefficients": relerr(port.coefficients(e["R"], e["Q"], A), d["coef_A"])}
        d.update(ext_W_cols=e["W"][:, cols] * fl, ext_P_cols=e["P"][:, cols] * fl, ext_R_cols=e["R"][:, cols] * fl,
                 ext_coef_A=port.coefficients(e["R"], e["Q"], A))
    print(tag, "reference vs extended:", dist["reference_type1_vs_extended"], f"({dist['ext_seconds']:.0f} s)", flush=True)
    d["distances_json"] = np.array(json.dumps(dist))
    save(tag, **d)
    json.dump(dist, open(os.path.join(HERE, tag + "_distances.json"), "w"), indent=1)


if __name__ == "__main__":
    assert ref.available(), "needs /root/reference (oracle/_ref/libplsref.so)"
    which = sys.argv[1:] or ["c5", "c4"]
    if "c5" in which:
        c5()
    if "c4" in which:
        c4()
    if "ext" in which or "c5" in which:
        ext("c5_full", C5, None)
    if "ext" in which or "c4" in which:
        ext("c4_full", C4, C4_TROWS)
