#!/usr/bin/env python
"""bench.py — headline benchmark of the PLS fit hot path (BASELINE.json: fits/sec + pass GB/s vs peak).

Workload (config 4 of BASELINE.json): synthetic X 1,000,000 x 2048, Y x 8, FP64, 32 components,
KERNEL_TYPE1 single fit, rows sharded over the ranks (strong scaling: the total work is fixed).
One "step" = one complete fit (X'Y pass + 32 fused passes + 33 single-CTA steps).

  python bench.py [--gpus N --steps K --warmup W]          our arm (CUDA, through the C-ABI)
  python bench.py --impl reference ...                      the reference's own CPU code (oracle/_ref)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C4 = dict(N=1_000_000, K=2048, M=8, A=32, L=16, seed=20260004)
C5 = dict(N=20_000, K=4096, M=4, A=64, L=16, seed=20260005)
WORKLOAD = "C4: synthetic X 1000000x2048, Y x8, FP64, 32 components, KERNEL_TYPE1 single fit"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.proc = index, [], False, None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")] + [time.perf_counter()])
                if self.stop_flag:
                    break
        except Exception:
            pass

    def finish(self, since=None):
        """Samples taken at or after `since` (perf_counter; the start of the timed region).  The sampler is started
        before the warm-up fits because nvidia-smi needs a few hundred ms to come up — longer than the timed region
        of a multi-GPU run; if the region was too short to catch two samples, the warm-up's (same load) are used."""
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:      # a sampler that lingers would keep polling the driver during the later legs
                self.proc.kill()
        if since is not None:
            inside = [r for r in self.rows if r[-1] >= since]
            if len(inside) >= 2:
                self.rows = inside
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        mx = next((int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()), None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_fit(n_sample, steps, warmup):
    """Time the reference's own plsr (oracle/_ref/libplsref.so) — else the C port — on the first
    n_sample rows of the C4 data; returns (seconds per fit on the sample, kind, cores)."""
    from oracle import port, ref
    X, Y = port.synth(C4["seed"], n_sample, C4["K"], C4["M"], L=C4["L"])
    if ref.available():
        for _ in range(warmup):
            ref.time_fit(X[:2000], Y[:2000], C4["A"], 0, 1)
        t = [ref.time_fit(X, Y, C4["A"], 0, 1) for _ in range(steps)]
        return sum(t) / len(t), "reference", 1
    t0 = time.time()
    for _ in range(steps):
        port.fit(X, Y, C4["A"], 0, want_T=True)
    return (time.time() - t0) / steps, "port", 1


def cpu_reference_cv(n_sample):
    """The reference's plsr per LOO fold (inner A = 64, the first A components of its own cv_LOO are bit-identical),
    one fold per host thread over all cores, on the first n_sample rows of the C5 matrix; returns
    (fits/s scaled to N = 20000, threads, seconds)."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    from oracle import port, ref
    if not ref.available():
        return None
    X, Y = port.synth(C5["seed"], n_sample, C5["K"], C5["M"], L=C5["L"])
    nthr = os.cpu_count() or 1
    te = np.arange(nthr, dtype=np.int64).reshape(-1, 1)
    t0 = time.time()
    with ThreadPoolExecutor(nthr) as ex:     # ctypes releases the GIL: the folds run concurrently
        list(ex.map(lambda f: ref.cv_folds(X, Y, C5["A"], 0, te[f:f + 1]), range(nthr)))
    sec = time.time() - t0
    return nthr / (sec * C5["N"] / n_sample), nthr, sec


def small_cv_legs(ctx, with_cpu=True, as_built=False):
    """C2 (nir/octane 60 x 401, A = 10, LOO: 60 folds) and C3 (PLS2 60 x 399, Y x3, A = 10, LSO: 1000 folds of 18 held-out
    rows) from the committed fixtures (tests/golden/, generated from the reference's own CSV files): fits/s through
    plsb_cv_folds with host index lists in and the residual block back on the host, PRESS checked against the fixture."""
    import numpy as np
    from pls_b200 import _capi
    out = {}
    gdir = os.path.join(ROOT, "tests", "golden")
    for name, key in (("c2_nir_octane", "loo"), ("c3_nir_pls2", "lso")):
        g = np.load(os.path.join(gdir, name + ".npz"))
        X, Y = np.asfortranarray(g["X"]), np.asfortranarray(g["Y"])
        te = np.arange(X.shape[0], dtype=np.int64).reshape(-1, 1) if key == "loo" else np.ascontiguousarray(g["lso_test"], dtype=np.int64)
        ds = _capi.Dataset(ctx, X, Y)
        E = ds.cv_folds(10, 0, te)
        reps = 20
        ctx.sync(); t0 = time.perf_counter()
        for _ in range(reps):
            E = ds.cv_folds(10, 0, te)                 # a fresh residual block per call (what Model::cv_* returns)
        ctx.sync(); sec_fresh = (time.perf_counter() - t0) / reps
        Eout = np.zeros((Y.shape[1], 10, te.size))
        ds.cv_folds(10, 0, te, out=Eout)
        ctx.sync(); t0 = time.perf_counter()
        for _ in range(reps):
            E = ds.cv_folds(10, 0, te, out=Eout)       # the caller's block reused (no first-touch page faults in the copy)
        ctx.sync(); sec = (time.perf_counter() - t0) / reps
        ctx.sync(); t0 = time.perf_counter()
        for _ in range(reps):
            ds.cv_press(10, 0, te)                     # validation's sums only: the block never leaves the device
        ctx.sync(); sec_press = (time.perf_counter() - t0) / reps
        press = (E.reshape(Y.shape[1], -1, 10) ** 2).sum(axis=1)
        want = g[key + "_ress"] if (key + "_ress") in g.files else None
        err = float(np.abs(press - want).max() / np.abs(want).max()) if want is not None else None
        out[name] = {"folds": int(te.shape[0]), "held_out_rows": int(te.shape[1]), "fits_per_s": te.shape[0] / sec,
                     "ms_per_call": sec * 1e3, "ms_per_call_fresh_output_block": sec_fresh * 1e3,
                     "ms_per_call_press_only": sec_press * 1e3, "press_rel_err_vs_reference": err}
        ds.close()
        out[name]["_data"] = (X, Y, te, g, key, E)
    for name in list(out):
        X, Y, te, g, key, E = out[name].pop("_data")
        if with_cpu:
            # the reference's plsr per fold (inner A = 10; its own cv_LOO/cv_LSO extract K components per fold and
            # are ~40x slower for the same first 10), folds split over all host threads
            try:
                from concurrent.futures import ThreadPoolExecutor
                from oracle import ref
                if ref.available():
                    nthr = os.cpu_count() or 1
                    nf = te.shape[0]
                    tr = np.ascontiguousarray(g["lso_train"], dtype=np.int64) if key == "lso" else None
                    cuts = [nf * i // nthr for i in range(nthr + 1)]
                    jobs = [(cuts[i], cuts[i + 1]) for i in range(nthr) if cuts[i + 1] > cuts[i]]
                    t0 = time.perf_counter()
                    with ThreadPoolExecutor(len(jobs)) as ex:
                        list(ex.map(lambda ab: ref.cv_folds(X, Y, 10, 0, te[ab[0]:ab[1]],
                                                            None if tr is None else tr[ab[0]:ab[1]]), jobs))
                    csec = time.perf_counter() - t0
                    out[name]["cpu_baseline"] = {"value": nf / csec, "unit": "fits/s", "cores": len(jobs), "kind": "reference",
                                                 "sample": f"all {nf} folds, reference plsr per fold (inner A = 10), {csec:.2f} s"}
                    if as_built and key == "loo":
                        # the reference's cv_LOO exactly as written: every fold extracts K = 401 components (pls.cpp:477)
                        t0 = time.perf_counter()
                        Er = ref.Model(X, Y, 10, 0).cv_loo()
                        asec = time.perf_counter() - t0
                        out[name]["cpu_as_built"] = {"value": nf / asec, "unit": "fits/s", "cores": 1, "kind": "reference",
                                                     "sample": f"Model::cv_LOO() unmodified, {asec:.1f} s",
                                                     "max_abs_diff_vs_gpu": float(np.abs(Er - E).max())}
            except Exception as e:
                out[name]["cpu_baseline"] = {"error": repr(e)}
    return out


def synth_host_parallel(c, rows):
    """The C4 / C5 generator on all host threads (row blocks are independent; ctypes releases the GIL)."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    from oracle import port
    K, M = c["K"], c["M"]
    X = np.zeros((rows, K), order="F"); Y = np.zeros((rows, M), order="F")
    nthr = os.cpu_count() or 1
    cuts = [rows * i // (4 * nthr) for i in range(4 * nthr + 1)]
    lib = port.lib()
    import ctypes as C

    def job(i):
        a, b = cuts[i], cuts[i + 1]
        if b > a:
            lib.orc_synth(C.c_uint64(c["seed"]), port.i64(a), port.i64(b - a), port.i64(K), port.i64(M), port.i64(c["L"]),
                          C.cast(X.ctypes.data + 8 * a, C.POINTER(C.c_double)), port.i64(rows),
                          C.cast(Y.ctypes.data + 8 * a, C.POINTER(C.c_double)), port.i64(rows))
    with ThreadPoolExecutor(nthr) as ex:
        list(ex.map(job, range(4 * nthr)))
    return X, Y


def run_reference(args):
    """The reference's own CPU implementation of the path on this box's host cores, on OUR arm's configuration:
    Model(X, Y, KERNEL_TYPE1, 32) on the full 1,000,000 x 2048 C4 matrix, unmodified (oracle/_ref/libplsref.so).  The
    reference is single-threaded as built, a full-size fit takes minutes, so ONE full-size fit is timed whatever --steps
    says (the line's `steps` is the number of fits actually timed); a few sampled fits (first 30000 rows, scaled) are
    kept beside it as a variance estimate, and the -fopenmp build of the same sources (all cores; Eigen threads its
    GEMM only) is timed on a 200000-row sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import port, ref
    N, K, M, A = C4["N"], C4["K"], C4["M"], C4["A"]
    full = args.ref_rows >= N
    n_gen = N if full else max(args.ref_rows, 2000)
    t0 = time.time()
    X, Y = synth_host_parallel(C4, n_gen)
    gen_s = time.time() - t0
    kind = "reference" if ref.available() else "port"

    def fit_seconds(rows):
        if kind == "reference":
            return ref.time_fit(X[:rows], Y[:rows], A, 0, 1)
        t = time.time(); port.fit(X[:rows], Y[:rows], A, 0, want_T=True); return time.time() - t
    fit_seconds(2000)                                             # warm-up (library load, page cache)
    n_s = min(30000, n_gen)
    sampled = [fit_seconds(n_s) * N / n_s for _ in range(min(max(args.steps - 1, 1), 3))]
    if full:
        sec_full, steps_timed, rows_timed = fit_seconds(N), 1, N
        sample = (f"ONE full-size fit: the unmodified reference's Model(X, Y, KERNEL_TYPE1, 32) on all {N} rows of the C4 matrix, "
                  f"single thread (the reference has no threading), {sec_full:.1f} s")
    else:
        sec_full, steps_timed, rows_timed = sum(sampled) / len(sampled), len(sampled), n_s
        sample = f"reference plsr on the first {n_s} rows, scaled by N/{n_s} (--ref-rows {args.ref_rows}: NOT the full-size fit)"
    omp = None
    try:
        n_o = min(200000, n_gen)
        t1 = ref.time_fit_omp(X[:n_o], Y[:n_o], A, 0, 1)
        if t1 is not None:
            omp = {"cores": os.cpu_count(), "type1_fits_per_s": 1.0 / (t1 * N / n_o),
                   "sample": f"-fopenmp build of the same sources, first {n_o} rows, {t1:.1f} s, scaled by N/{n_o}"}
    except Exception as e:
        omp = {"error": repr(e)}
    val = 1.0 / sec_full
    line = {"impl": "reference", "metric": "pls_fits_per_sec", "value": val, "unit": "fits/s", "n_gpus": args.gpus,
            "steps": steps_timed, "steps_requested": args.steps, "warmup": 1, "ms_per_step": sec_full * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD},
            "config_notes": {"sample_rows": rows_timed, "parallelism": "1 host thread (the reference has no threading)"},
            "cpu_baseline": {"value": val, "unit": "fits/s", "cores": 1, "kind": kind, "sample": sample,
                             "sampled_fits_per_s_scaled": [1.0 / x for x in sampled], "openmp_build": omp,
                             "input_generation_s": gen_s},
            "e2e": {"value": val, "unit": "fits/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def cpp_dropin_legs(c4, c5, device, with_cv=True, timeout=900):
    """tests/cpp/api_bench: the C++ drop-in API on pageable Eigen matrices, timed with the host clock inside the program."""
    exe = os.path.join(ROOT, "tests", "cpp", "api_bench")
    if not os.path.exists(exe):
        return {"error": "tests/cpp/api_bench not built (needs Eigen headers at build time)"}
    env = dict(os.environ, PLSB_DEVICES=str(device))
    out = {}
    jobs = [("fit_c4", ["fit", c4["N"], c4["K"], c4["M"], c4["A"], 3, c4["seed"], c4["L"]])]
    if with_cv:
        jobs.append(("loo_c5", ["loo", c5["N"], c5["K"], c5["M"], c5["A"], 1, c5["seed"], c5["L"]]))
    for name, argv in jobs:
        try:
            r = subprocess.run([exe] + [str(a) for a in argv], capture_output=True, text=True, timeout=timeout, env=env)
            out[name] = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else {"error": r.stderr[-500:]}
        except Exception as e:
            out[name] = {"error": repr(e)}
    out["what"] = ("PLS::Model(X, Y, KERNEL_TYPE1, A) constructed from an ordinary Eigen::MatrixXd (pageable host memory) and "
                   "cv_LOO() + validation(MSE) + optimal_num_components(), seconds per call by the host clock in C++")
    return out


def parity_vs_reference_fixture(ds, A, r0, r1, dist, torch):
    """One more fit of the timed configuration with its results read back, compared with the reference's own fit of the
    same matrix (committed fixture).  Returns the worst relative errors (max over ranks); raises if any exceeds the
    north star's 1e-9 (W: 4.5e-9, 2.5 x the distance between the reference's own two kernels — tests/test_gpu_fullsize.py)."""
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", "c4_full.npz")
    if not os.path.exists(path):
        return {"error": "tests/golden/c4_full.npz missing"}
    g = np.load(path)
    st = ds.fit_resident(A, 0, download="WPRQT")

    def signs(R):
        sg = np.sign(R[np.abs(R).argmax(axis=0), np.arange(R.shape[1])]); sg[sg == 0] = 1.0
        return sg

    def col_err(a, b):
        den = np.abs(b).max(axis=0); den[den == 0] = 1.0
        return float((np.abs(a - b).max(axis=0) / den).max())

    sg = signs(st["R"]) * signs(g["R"])
    errs = {k: col_err(st[k] * sg, g[k]) for k in "WPRQ"}
    B = st["R"] @ st["Q"].T
    errs["coefficients"] = float(np.abs(B - g["coef_A"]).max() / np.abs(g["coef_A"]).max())
    rows = g["T_rows"]; mine = (rows >= r0) & (rows < r1)
    errs["T"] = col_err(st["T"][rows[mine] - r0] * sg, g["T_sub"][mine]) if mine.any() else 0.0
    errs["coef_checksum"] = float(np.abs(B).sum())
    keys = sorted(errs)
    if dist is not None:
        t = torch.tensor([errs[k] for k in keys], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        errs = dict(zip(keys, t.tolist()))
    # W: 2.5 x the distance between the reference's own two kernels, read from the fixture (tests/test_gpu_fullsize.py)
    w_bound = 2.5 * json.loads(str(g["distances_json"]))["reference_type2_vs_reference_type1"]["W"]
    bad = {k: v for k, v in errs.items() if k != "coef_checksum" and v > (w_bound if k == "W" else 1e-9)}
    if bad:
        raise SystemExit(f"bench.py: the timed configuration does not reproduce the reference's fit: {bad}")
    errs["W_bound"] = w_bound
    errs["what"] = ("worst relative error (per component after sign normalisation; coefficients matrix-relative) of this "
                    "run's fit vs the unmodified reference's KERNEL_TYPE1 fit of the same 1M x 2048 matrix; max over ranks")
    return errs


def cv_parity_vs_reference_fixture(E, f0, f1, dist, torch):
    """Fold-sharded LOO at C5 against the UNMODIFIED reference (pls.cpp:469-491): tests/golden/c5_full.npz holds the
    residuals of the 8 folds the reference refitted on the full 20000 x 4096 matrix.  Each rank checks the fixture folds
    that fall in its shard [f0, f1) of the residual block E[m][c][row] the timed call filled; worst matrix-relative error,
    max over ranks, and the number of folds checked, summed over ranks (must be all 8).  Raises above 1e-9."""
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", "c5_full.npz")
    if not os.path.exists(path):
        return {"error": "tests/golden/c5_full.npz missing"}
    g = np.load(path)
    folds, want = g["loo_folds"], g["loo"]                    # (8,), [M][8][A]
    err, n = 0.0, 0
    for j, i in enumerate(folds):
        if f0 <= i < f1:
            err = max(err, float(np.abs(E[:, :, i] - want[:, j, :]).max() / np.abs(want[:, j, :]).max()))
            n += 1
    if dist is not None:
        t = torch.tensor([err], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        c = torch.tensor([n], device="cuda", dtype=torch.int64); dist.all_reduce(c, op=dist.ReduceOp.SUM)
        err, n = t.item(), int(c.item())
    if err > 1e-9 or n != len(folds):
        raise SystemExit(f"bench.py: fold-sharded LOO at C5 does not reproduce the reference's refits: worst error {err:.3e} "
                         f"over {n} of {len(folds)} fixture folds")
    return {"rel_err": err, "folds_checked": n,
            "what": "residuals (all responses, all 64 component counts) of the folds the unmodified reference refitted, "
                    "each checked by the rank whose shard holds it; matrix-relative, max over ranks"}


def loo_c4_parity(ctx, ds4, te4, press, g0, g1, A, rank, world, dist, torch, write_fixture):
    """cv_LOO on the 1M x 2048 matrix, checked two ways.  (1) One fold of this rank's shard: the engine's residuals for
    a 640-fold slice around it (two lanes, like the timed call) against an EXPLICIT refit on the other 999,999 rows by the
    streaming-pass fit — the path `parity` pins on the reference's own fit of this matrix — and prediction of the
    left-out row (pls.cpp:478-489).  (2) The PRESS matrix summed over all ranks against tests/golden/c4_loo_press.json,
    the value of a single-GPU run whose folds were spot-checked the same way.  Raises above 1e-9."""
    import numpy as np
    from pls_b200 import _capi
    N, K, M = ds4.N, ds4.K, ds4.M
    i = g0 + (g1 - g0) // 2
    lo = max(g0, min(i - 320, g1 - 640))
    te_s = te4[lo:lo + 640]
    E = ds4.cv_folds(A, 0, te_s)                                             # [M][640][A]
    parts = [_capi.Dataset(ctx, synth=(C4["seed"], 0, i, K, M, C4["L"])),
             _capi.Dataset(ctx, synth=(C4["seed"], i + 1, N - i - 1, K, M, C4["L"]))]
    rest = _capi.Dataset.concat(ctx, parts)
    [p.close() for p in parts]
    st = rest.fit_resident(A, 0, download="RQ")
    rest.close()
    one = _capi.Dataset(ctx, synth=(C4["seed"], i, 1, K, M, C4["L"]))
    Xi, Yi = one.download()
    one.close()
    res = _capi.predict_residuals(ctx, st["R"], st["Q"], Xi, Yi)             # [M][1][A]
    e_fold = float(np.abs(E[:, i - lo, :] - res[:, 0, :]).max() / np.abs(res[:, 0, :]).max())
    ctx.trim()
    path = os.path.join(ROOT, "tests", "golden", "c4_loo_press.json")
    e_press = None
    if write_fixture and rank == 0 and world == 1:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump({"what": "PRESS[m][c] of cv_LOO on the C4 matrix (1M folds, 32 components), single-GPU run of bench.py "
                           "--write-fixtures; copy to tests/golden/c4_loo_press.json",
                   "press": [[float(v) for v in row] for row in press]},
                  open(os.path.join(ROOT, "gpurun_out", "c4_loo_press.json"), "w"))
    if os.path.exists(path):
        want = np.array(json.load(open(path))["press"])
        e_press = float((np.abs(press - want) / np.abs(want)).max())
    if dist is not None:
        t = torch.tensor([e_fold], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e_fold = t.item()
    if e_fold > 1e-9 or (e_press is not None and e_press > 1e-9):
        raise SystemExit(f"bench.py: cv_LOO on the C4 matrix fails parity: explicit refit of fold {i} {e_fold:.3e}, "
                         f"PRESS vs fixture {e_press}")
    return {"explicit_refit_rel_err": e_fold, "explicit_refit_fold_rank0": int(i), "press_vs_fixture_rel_err": e_press,
            "what": "one fold per rank refitted explicitly on the other 999,999 rows (max over ranks); PRESS summed over "
                    "ranks vs the committed single-GPU value, element-wise relative"}


def run_ours(args):
    import numpy as np
    import torch
    from pls_b200 import _capi
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = _capi.Context(local)
    # nvidia-smi needs up to a second to come up when several ranks start one each: start sampling now, keep only the
    # samples of the timed region (ClockSampler.finish)
    sampler = ClockSampler(local)
    if rank == 0:      # one poller per job: the line rank 0 prints carries its GPU's clocks
        sampler.start()
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(_capi.Context.unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        ctx.comm_init(rank, world, bytes(uid.cpu().numpy().tobytes()))

    N, K, M, A = C4["N"], C4["K"], C4["M"], C4["A"]
    from pls_b200 import sharding
    r0, r1 = sharding.row_shard(N, rank, world)
    ds = _capi.Dataset(ctx, synth=(C4["seed"], r0, r1 - r0, K, M, C4["L"]), sharded=world > 1)

    def barrier():
        ctx.sync()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize() if torch.cuda.is_available() else None

    # ---- device-resident timing: K steps bracketed by barrier + sync; per-kernel CUDA events inside ----
    # Every timed fit is the COMPLETE KERNEL_TYPE1 fit of the reference (pls.cpp:390-437): W, P, R, Q and the N x A
    # scores T are formed and stored in HBM (plsb_fit_resident); nothing is read back inside the timed region.
    for _ in range(max(args.warmup, 3)):
        ds.fit_resident(A, 0)
    # (1) the timed region: exactly K fits; one CUDA event pair per fit on the library's stream (device time), no
    # event between the kernels of a fit
    ctx.profile(2); ctx.profile_reset()
    barrier()
    l0 = ctx.launch_count()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ds.fit_resident(A, 0)
    barrier()
    wall = time.perf_counter() - t0
    launches = ctx.launch_count() - l0
    dev_timed_ms, n_timed = ctx.profile_get("fit_total")
    # (2) the same K fits again with a CUDA event pair around every launch (library stream): per-kernel durations
    # for the roofline.  Kept out of (1) because an event record between two kernels switches off the programmatic
    # dependent launch that lets a kernel's set-up overlap its predecessor's tail.
    ctx.profile(True); ctx.profile_reset()
    barrier()
    t1 = time.perf_counter()
    for _ in range(args.steps):
        ds.fit_resident(A, 0)
    barrier()
    wall_instr = time.perf_counter() - t1
    clocks = sampler.finish(since=t0) if rank == 0 else None
    dev_ms, nfit = ctx.profile_get("fit_total")
    pass_ms, npass = ctx.profile_get("pass_fused")
    xty_ms, nxty = ctx.profile_get("pass_xty")
    step_ms, nstep = ctx.profile_get("step")
    red_ms, nred = ctx.profile_get("reduce")
    ctx.profile(False)
    # device time of the K fits (CUDA events) and the host clock around the bracketed region: report the larger
    ms_per_step = max(wall * 1e3 / args.steps, dev_timed_ms / max(n_timed, 1))
    if dist is not None:
        t = torch.tensor([ms_per_step, wall_instr * 1e3 / args.steps, pass_ms / max(npass, 1)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_per_step, instr_ms, pass_avg = t.tolist()
    else:
        instr_ms, pass_avg = wall_instr * 1e3 / args.steps, pass_ms / max(npass, 1)
    value = 1e3 / ms_per_step

    # ---- parity of exactly this configuration (same data, same sharding, same kernels) against the UNMODIFIED
    # reference's fit of the C4 matrix (tests/golden/c4_full.npz: its W, P, R, Q, coefficients and 192 rows of T) ----
    parity = parity_vs_reference_fixture(ds, A, r0, r1, dist, torch) if not args.no_parity else None
    comm = ctx.comm_info()

    # ---- roofline of the dominant kernel (fused pass): algorithmic bytes = 8*N*K read + 8*N t written -----
    rows = r1 - r0
    alg_bytes = 8.0 * rows * K + 8.0 * rows
    peak, which = peaks()
    achieved = alg_bytes / (pass_avg * 1e-3) / 1e9
    traffic = None
    tfile = os.path.join(ROOT, "profiles", "traffic_fused_pass.json")
    if os.path.exists(tfile) and world == 1:
        try:
            traffic = json.load(open(tfile)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": "fused_pass_kernel (g = X'(X r), tt, t: one read of X)", "achieved": achieved,
                "peak": peak, "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({which})", "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic,
                "traffic_source": ("profiles/traffic_fused_pass.json: dram__bytes_read.sum + dram__bytes_write.sum of one "
                                   "`ncu --set full` capture of this kernel on the N = 1 workload (a committed constant, "
                                   "not measured in this run)") if traffic is not None else
                                  "none: the ncu capture is of the full 1M-row launch, this rank's shard is smaller",
                "algorithmic_bytes_per_launch": alg_bytes,
                "avg_launch_ms": pass_avg, "launches_timed": npass,
                "other_kernels_ms_per_fit": {"xty_pass": xty_ms / max(nfit, 1), "steps": step_ms / max(nfit, 1),
                                             "partial_reductions": red_ms / max(nfit, 1)}}

    # ---- the one dense contraction of a KERNEL_TYPE2 fit at C4: XX = X'X (pls.cpp:398) on the FP64 tensor cores -----
    xtx = None
    if not args.no_xtx:
        ds.xtx(download=False)                                   # warm-up
        ctx.profile(True); ctx.profile_reset()
        ds.xtx(download=False)
        xtx_ms, _ = ctx.profile_get("xtx")
        ctx.profile(False)
        if dist is not None:
            t = torch.tensor([xtx_ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            xtx_ms = t.item()
        issued = float(N) * K * (K + 128)                        # SYRK: upper-triangular 128-wide tiles only
        xtx = {"workload": "XX = X'X of the C4 matrix (rows sharded), incl. the cross-rank sum", "ms": xtx_ms,
               "tflops_as_reference_counts_2NK2": 2.0 * N * K * K / (xtx_ms * 1e-3) / 1e12,
               "tflops_issued_syrk": issued / (xtx_ms * 1e-3) / 1e12}

    # ---- end to end through the public API with HOST buffers (upload + fit + results back) --------------
    e2e = None
    if not args.no_e2e:
        Xh_t = torch.empty((K, rows), dtype=torch.float64).pin_memory()
        Xh = Xh_t.numpy().T                                                          # column-major rows x K
        Yh = torch.empty((M, rows), dtype=torch.float64).pin_memory().numpy().T
        _capi.check(_capi.lib().plsb_dataset_download(ds.h, _capi.P(Xh), _capi.i64(rows), _capi.P(Yh), _capi.i64(rows)))
        ds.close()
        Wt = np.zeros((K, A), order="F"); Pt = np.zeros((K, A), order="F"); Rt = np.zeros((K, A), order="F")
        Qt = np.zeros((M, A), order="F"); Tt = torch.empty((A, rows), dtype=torch.float64).pin_memory().numpy().T

        def one():
            dsh = _capi.Dataset(ctx, Xh, Yh, sharded=world > 1)
            _capi.check(_capi.lib().plsb_fit(dsh.h, _capi.i64(A), 0, _capi.P(Wt), _capi.P(Pt), _capi.P(Rt), _capi.P(Qt),
                                             _capi.P(Tt)))
            dsh.close()
        one()
        ctx.profile(True); ctx.profile_reset()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            one()
        barrier()
        e2e_s = (time.perf_counter() - t0) / args.e2e_steps
        up_ms, nup = ctx.profile_get("upload"); fit_ms, nf = ctx.profile_get("fit_total")
        sc_ms, _ = ctx.profile_get("scores"); pf_ms, npf = ctx.profile_get("pass_fused")
        ctx.profile(False)
        if dist is not None:
            t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_s = t.item()
        e2e = {"value": 1.0 / e2e_s, "unit": "fits/s", "h2d_bytes_per_step": int(8 * rows * (K + M)),
               "d2h_bytes_per_step": int(8 * (3 * K * A + M * A + rows * A)), "ms_per_step": e2e_s * 1e3,
               "steps": args.e2e_steps, "what": "plsb_dataset_create(pinned host X,Y) + plsb_fit + W,P,R,Q,T to host",
               "h2d_GBps_rank0": 8.0 * rows * (K + M) / (up_ms / max(args.e2e_steps, 1) * 1e-3) / 1e9 if up_ms > 0 else None,
               "device_ms_rank0": {"upload_incl_xtx_xty_under_the_transfer": up_ms / max(args.e2e_steps, 1),
                                   "fit_after_upload": fit_ms / max(nf, 1), "of_which_scores_T": sc_ms / max(nf, 1)},
               "path": ("X'X and X'Y accumulated on the tensor cores while the rows cross PCIe; components from the "
                        "L2-resident X'X; one DMMA pass T = X R") if npf == 0 else "streaming passes (one read of X per component)"}

        # the denominator of the end-to-end figure: what this box's PCIe / host memory gives when every rank copies its
        # pinned shard to its GPU at the same time and nothing else runs (no kernels, no packing, no results back)
        try:
            dbuf = torch.empty((K, rows), dtype=torch.float64, device="cuda")
            dbuf.copy_(Xh_t, non_blocking=True); torch.cuda.synchronize()
            barrier()
            t0 = time.perf_counter()
            for _ in range(2):
                dbuf.copy_(Xh_t, non_blocking=True)
            torch.cuda.synchronize()
            barrier()
            copy_s = (time.perf_counter() - t0) / 2
            del dbuf
            if dist is not None:
                t = torch.tensor([copy_s], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                copy_s = t.item()
            floor_s = copy_s * (K + M) / K                        # X and Y
            e2e["h2d_floor_ms"] = floor_s * 1e3
            e2e["h2d_floor_GBps_per_rank"] = 8.0 * rows * K / copy_s / 1e9
            e2e["frac_of_h2d_floor"] = floor_s / e2e_s
            e2e["h2d_floor_what"] = (f"{world} concurrent cudaMemcpyAsync of each rank's pinned X shard ({8.0 * rows * K / 1e9:.2f} GB), "
                                     "max over ranks, scaled to X and Y; frac_of_h2d_floor = floor / measured e2e step")
        except Exception as e:
            e2e["h2d_floor_ms"] = None
            e2e["h2d_floor_what"] = repr(e)
        del Xh_t

    # ---- config 5: full leave-one-out CV, 20 000 fits of 20000 x 4096 (A = 64), folds sharded over the ranks -----
    cv = None
    if not args.no_cv:
        if args.no_e2e:
            ds.close()      # (the e2e leg released the C4 data set itself)
        ctx.trim()
        n5, k5, m5, a5 = C5["N"], C5["K"], C5["M"], C5["A"]
        ds5 = _capi.Dataset(ctx, synth=(C5["seed"], 0, n5, k5, m5, C5["L"]))      # replicated on every rank
        te = np.arange(n5, dtype=np.int64).reshape(-1, 1)
        f0, f1 = sharding.fold_shard(n5, rank, world)
        E = np.zeros((m5, a5, n5))
        ds5.cv_folds(a5, 0, te, fold0=f0, nfolds_local=f1 - f0, out=E)   # warm-up: the whole call (buffers, both lanes)
        # (1) the timed call: X'X and X'Y once (each rank reads its row slice, partial matrices all-reduced), then the
        # folds of this rank — two lanes in flight, no event between kernels
        ctx.profile(3); ctx.profile_reset()
        barrier()
        t0 = time.perf_counter()
        ds5.build_gram(across_ranks=world > 1)
        xtx_ms, _ = ctx.profile_get("xtx")
        ctx.profile(False)
        ds5.cv_folds(a5, 0, te, fold0=f0, nfolds_local=f1 - f0, out=E)              # includes D2H of the residual block
        barrier()
        cv_s = time.perf_counter() - t0
        # (2) the same call with an event pair around every kernel, lanes kept: kernel times INSIDE the overlap
        ctx.profile(3); ctx.profile_reset()
        t1 = time.perf_counter()
        ds5.cv_folds(a5, 0, te, fold0=f0, nfolds_local=f1 - f0, out=E)
        barrier()
        cv_instr_s = time.perf_counter() - t1
        gemm_ov, ngemm = ctx.profile_get("cv_gemm"); fold_ov, _ = ctx.profile_get("cv_fold")
        # (3) the same kernels and buffers one after the other (PLSB_CV_SERIAL): each kernel alone on the GPU — the
        # product kernel's FP64 tensor roofline, the fold kernels' HBM time
        ctx.profile_reset()
        os.environ["PLSB_CV_SERIAL"] = "1"
        try:
            t2 = time.perf_counter()
            ds5.cv_folds(a5, 0, te, fold0=f0, nfolds_local=f1 - f0, out=E)
            barrier()
            cv_serial_s = time.perf_counter() - t2
        finally:
            del os.environ["PLSB_CV_SERIAL"]
        gemm_ms, _ = ctx.profile_get("cv_gemm"); fold_ms, _ = ctx.profile_get("cv_fold")
        ctx.profile(False)
        if dist is not None:
            t = torch.tensor([cv_s], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            cv_s = t.item()
        gemm_flops = 2.0 * k5 * k5 * a5 * (f1 - f0)
        fp64_peak = None
        try:    # calibration only (never on the product path): cuBLAS DGEMM 8192^3 as the FP64 tensor denominator
            a = torch.randn(8192, 8192, device="cuda", dtype=torch.float64); b = torch.randn_like(a)
            torch.matmul(a, b); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
            fp64_peak = 2 * 8192 ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12
            del a, b
        except Exception:
            pass
        gemm_tf = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else None
        cv = {"workload": "C5: synthetic X 20000x4096, Y x4, FP64, 64 components, full LOO-CV (20000 fits), folds sharded",
              "fits_per_s": n5 / cv_s, "seconds": cv_s, "folds": n5, "folds_this_rank": f1 - f0,
              "schedule": "two lanes: the persistent DMMA product of one wave under the HBM-bound fold kernel of the other",
              "kernels_ms_rank0": {"alone": {"dmma_gemm": gemm_ms, "fold_kernels": fold_ms, "call_s": cv_serial_s},
                                   "overlapped": {"dmma_gemm": gemm_ov, "fold_kernels": fold_ov, "call_s": cv_instr_s},
                                   "xtx": xtx_ms},
              "gemm_roofline": {"bound": "tensor", "kernel": "cv_gemm_kernel, timed alone", "achieved": gemm_tf, "peak": fp64_peak,
                                "unit": "TFLOP/s", "frac": (gemm_tf / fp64_peak) if (gemm_tf and fp64_peak) else None,
                                "peak_source": "cuBLAS DGEMM 8192^3 via torch.matmul, measured in this run (calibration only)"}}
        cv["parity"] = cv_parity_vs_reference_fixture(E, f0, f1, dist, torch) if not args.no_parity else None
        ds5.close()

    # ---- north_star's target sentence: LOO-CV on the 1M x 2048 matrix itself (pls.cpp:469-491 at N = 10^6): one million
    # fits of A = 32 components, folds sharded over the ranks, every rank holding the whole matrix; X'X / X'Y built once
    # (each rank reads its row slice, partial matrices all-reduced) and PRESS reduced on the device (the residual block,
    # M*A*N*8 = 2 GB, never leaves HBM) -----
    loo_c4 = None
    if not args.no_cv and not args.no_loo_c4:
        try:
            ctx.trim()
            ds4 = _capi.Dataset(ctx, synth=(C4["seed"], 0, N, K, M, C4["L"]))
            te4 = np.arange(N, dtype=np.int64).reshape(-1, 1)
            g0, g1 = sharding.fold_shard(N, rank, world)
            ds4.cv_press(A, 0, te4, fold0=g0, nfolds_local=min(9472, g1 - g0))      # warm-up: one round of two waves
            ctx.profile(3); ctx.profile_reset()
            barrier()
            t0 = time.perf_counter()
            ds4.build_gram(across_ranks=world > 1)
            barrier()
            gram_s = time.perf_counter() - t0
            xtx4_ms, _ = ctx.profile_get("xtx"); xty4_ms, _ = ctx.profile_get("pass_xty")
            ctx.profile(False)
            barrier()
            t1 = time.perf_counter()
            press = ds4.cv_press(A, 0, te4, fold0=g0, nfolds_local=g1 - g0)
            barrier()
            loo_s = time.perf_counter() - t1
            if dist is not None:
                t = torch.tensor([gram_s, loo_s], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                gram_s, loo_s = t.tolist()
                pt = torch.tensor(np.ascontiguousarray(press), device="cuda", dtype=torch.float64)
                dist.all_reduce(pt, op=dist.ReduceOp.SUM)
                press = pt.cpu().numpy()
            rows4 = (N // world) if world > 1 else N
            loo_c4 = {"workload": "cv_LOO on the C4 matrix: 1,000,000 folds of 1M x 2048, Y x8, 32 components, folds sharded, PRESS on device",
                      "fits_per_s": N / (gram_s + loo_s), "seconds": gram_s + loo_s, "folds": N, "folds_this_rank": g1 - g0,
                      "cross_products": {"seconds": gram_s, "xtx_ms_rank0": xtx4_ms, "xty_ms_rank0": xty4_ms,
                                         "rows_read_per_rank": rows4,
                                         "xty_pass_GBps_rank0": 8.0 * rows4 * (K + M) / (xty4_ms * 1e-3) / 1e9 if xty4_ms > 0 else None,
                                         "xtx_tflops_issued_rank0": float(rows4) * K * (K + 128) / (xtx4_ms * 1e-3) / 1e12 if xtx4_ms > 0 else None},
                      "folds_seconds": loo_s, "gemm_tflops_if_alone": 2.0 * K * K * A * (g1 - g0) / loo_s / 1e12,
                      "mse_last_component": [float(v) for v in (press[:, A - 1] / N)]}
            if not args.no_parity:
                loo_c4["parity"] = loo_c4_parity(ctx, ds4, te4, press, g0, g1, A, rank, world, dist, torch, args.write_fixtures)
            ds4.close()
        except Exception as e:
            loo_c4 = {"error": repr(e)}

    # ---- configs 2 and 3: small-data CV (60 x ~400 spectra), one CTA per fold, all folds of a call in one wave -----
    cv_small = None
    if rank == 0 and not args.no_cv:
        try:
            cv_small = small_cv_legs(ctx, with_cpu=(world == 1 and not args.no_cpu), as_built=args.ref_as_built)
        except Exception as e:
            cv_small = {"error": repr(e)}

    # ---- the drop-in itself, end to end: PLS::Model(X, Y, KERNEL_TYPE1, 32) from an ordinary (pageable) Eigen matrix and
    # cv_LOO() + validation + optimal_num_components at C5, wall clock inside a C++ program (tests/cpp/api_bench.cpp) ----
    e2e_cpp = None
    if rank == 0 and world == 1 and not args.no_e2e:
        ctx.trim()
        e2e_cpp = cpp_dropin_legs(C4, C5, local, with_cv=not args.no_cv)
        if e2e is not None and e2e_cpp and "fit_c4" in e2e_cpp and "model_ctor_s_mean" in e2e_cpp["fit_c4"]:
            e2e["cpp_model_ctor_fits_per_s"] = 1.0 / e2e_cpp["fit_c4"]["model_ctor_s_mean"]

    # the same drop-in on ALL the GPUs of the run: ONE C++ process, PLSB_DEVICES=0..N-1, rows sharded by the engine's own
    # threads + peer exchange (host/model.cpp) — what a C++ caller of PLS::Model gets on this box.  The torchrun ranks
    # give the GPUs' memory back and wait at a barrier while rank 0 runs it.
    if world > 1 and not args.no_e2e and args.e2e_cpp_multi:
        ctx.trim()
        torch.cuda.empty_cache()
        barrier()
        if rank == 0:
            e2e_cpp = cpp_dropin_legs(C4, C5, ",".join(str(i) for i in range(world)), with_cv=False, timeout=240)
            if e2e is not None and e2e_cpp and "fit_c4" in e2e_cpp and "model_ctor_s_mean" in e2e_cpp["fit_c4"]:
                e2e["cpp_model_ctor_fits_per_s"] = 1.0 / e2e_cpp["fit_c4"]["model_ctor_s_mean"]
        barrier()

    # CPU baselines come last: sixteen busy host threads right before a latency-bound GPU leg slow its launches down
    if cv is not None and rank == 0 and world == 1 and not args.no_cpu:
        try:
            r = cpu_reference_cv(4000)
            if r:
                cv["cpu_baseline"] = {"value": r[0], "unit": "fits/s", "cores": r[1], "kind": "reference",
                                      "sample": f"{r[1]} LOO folds, one per host thread, reference plsr (A=64) on the first 4000 "
                                                f"rows of the C5 matrix ({r[2]:.1f} s), scaled by N/4000"}
        except Exception as e:
            cv["cpu_baseline"] = {"error": repr(e)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            sec, kind, cores = cpu_reference_fit(args.cpu_sample, 1, 1)
            cpu = {"value": 1.0 / (sec * N / args.cpu_sample), "unit": "fits/s", "cores": cores, "kind": kind,
                   "sample": f"one KERNEL_TYPE1 fit (A=32) on the first {args.cpu_sample} rows of the C4 matrix "
                             f"({sec:.2f} s), scaled by N/{args.cpu_sample}"}
            try:    # the reference built with -fopenmp (Eigen threads GEMM only): all cores, a 20000-row sample
                from oracle import port, ref
                Xo, Yo = port.synth(C4["seed"], 20000, K, M, L=C4["L"])
                t1, t2 = ref.time_fit_omp(Xo, Yo, A, 0, 1), ref.time_fit_omp(Xo, Yo, A, 1, 1)
                if t1 is not None:
                    cpu["openmp_build"] = {"cores": os.cpu_count(), "type1_fits_per_s": 1.0 / (t1 * N / 20000),
                                           "type2_fits_per_s": 1.0 / (t2 * N / 20000),
                                           "sample": f"20000 rows, type 1 {t1:.2f} s, type 2 {t2:.2f} s, scaled by N/20000"}
            except Exception as e:
                cpu["openmp_build"] = {"error": repr(e)}
        except Exception as e:  # the oracle always exists; report rather than hide a failure
            cpu = {"error": repr(e)}

    if rank == 0:
        line = {"metric": "pls_fits_per_sec", "value": value, "unit": "fits/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": WORKLOAD},        # the same dict on both arms: the comparison is like for like
                "config_notes": {"parallelism": f"rows sharded over {world} rank(s)",
                                 "l2": "inputs (16.4 GB of X per fit pass) are larger than L2; no flush needed"},
                "roofline": roofline, "parity": parity, "exchange": comm["exchange"], "cpu_baseline": cpu, "e2e": e2e, "e2e_cpp": e2e_cpp, "cv": cv, "loo_c4": loo_c4, "cv_small": cv_small, "xtx_c4": xtx, "gpu_launches": int(launches), "clocks": clocks,
                "device_ms_per_step": dev_timed_ms / max(n_timed, 1),
                "instrumented": {"ms_per_step": instr_ms, "device_ms_per_fit": dev_ms / max(nfit, 1),
                                 "what": "the same K fits repeated with a CUDA event pair around every launch: source of the per-kernel times"}}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=120000)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-cv", action="store_true")
    ap.add_argument("--no-xtx", action="store_true")
    ap.add_argument("--no-loo-c4", action="store_true", help="skip the 1,000,000-fold LOO on the C4 matrix (~10 s on one GPU)")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--e2e-cpp-multi", action="store_true",
                    help="N > 1: also time the C++ drop-in (one process, PLSB_DEVICES=0..N-1) while the ranks wait at a barrier")
    ap.add_argument("--write-fixtures", action="store_true",
                    help="N = 1 only: write gpurun_out/c4_loo_press.json (the PRESS fixture the loo_c4 parity check reads)")
    ap.add_argument("--ref-rows", type=int, default=C4["N"],
                    help="reference arm: rows of the C4 matrix its timed fit runs on (default: all 1,000,000; fewer = a "
                         "sample scaled by N/rows, flagged in the line)")
    ap.add_argument("--ref-as-built", action="store_true", help="also time the reference's own cv_LOO on config 2 (~35 s)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
