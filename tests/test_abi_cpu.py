"""CPU-side checks of the drop-in boundary: libplsb.so loads, exports every symbol include/plsb.h
declares, fails loudly without a device, and its host-side statistics (validation, Wilcoxon,
optimal_num_components, LSO fold lists) reproduce the reference's golden vectors."""
import os
import re

import numpy as np
import pytest

from pls_b200 import _capi
from tests.util import RTOL, load_golden, relerr

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "plsb.h")).read()
    declared = set(re.findall(r"\b(plsb_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"plsb_ctx", "plsb_dataset"}
    assert declared, "no declarations parsed"
    lib = _capi.lib()
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, f"libplsb.so does not export: {missing}"
    assert set(_capi.SYMBOLS) == declared


def test_no_cpu_fallback_without_device():
    import ctypes as C
    n = C.c_int(-1)
    rc = _capi.lib().plsb_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(_capi.PlsbError):
        _capi.Context(0)


def test_host_stats_match_reference_goldens():
    s = load_golden("stats")
    assert np.allclose([_capi.normalcdf(z) for z in s["z"]], s["cdf"], rtol=0, atol=1e-15)
    assert np.allclose([_capi.wilcoxon(a, b) for a, b in zip(s["e1"], s["e2"])], s["wil"], rtol=0, atol=1e-14)
    g = load_golden("c2_nir_octane")
    assert relerr(_capi.validation(g["loo"], 0), g["loo_ress"]) <= RTOL
    assert relerr(_capi.validation(g["loo"], 1), g["loo_mse"]) <= RTOL
    assert np.array_equal(_capi.optimal_num_components(g["loo"]), g["loo_opt"])
    g = load_golden("c1_toy")
    assert np.array_equal(_capi.optimal_num_components(g["loo"]), g["loo_opt"])
    assert np.array_equal(_capi.optimal_num_components(g["lso"]), g["lso_opt"])
    assert relerr(_capi.validation(g["lso"], 1), g["lso_mse"]) <= RTOL


def test_summaries_of_a_large_block_are_threaded_and_unchanged():
    """validation / optimal_num_components split a large residual block over host threads by response; every response is
    still summed and ranked by one thread in the reference's order: the oracle's single-threaded answers, bit for bit."""
    from oracle import port
    rng = np.random.default_rng(11)
    M, rows, A = 5, 40000, 9                     # rows * A >= 2^18: the threaded path
    E = rng.standard_normal((M, rows, A)) * (1.0 + 0.3 * rng.standard_normal((M, 1, A)) ** 2)
    assert np.array_equal(_capi.validation(E, 0), port.validation(E, 0))
    assert np.array_equal(_capi.validation(E, 1), port.validation(E, 1))
    assert np.array_equal(_capi.optimal_num_components(E), port.optimal_num_components(E))


@pytest.mark.parametrize("name,N,trials", [("c1_toy", 10, 100), ("c2_nir_octane", 60, 600), ("c3_nir_pls2", 60, 1000)])
def test_lso_folds_match_reference(name, N, trials):
    g = load_golden(name)
    tr, te = _capi.lso_folds(N, 0.3, trials)          # std::mt19937 default seed, like main.cpp:38
    assert np.array_equal(tr, g["lso_train"].astype(np.int64))
    assert np.array_equal(te, g["lso_test"].astype(np.int64))


# ---- the C++ drop-in layer's host-only behaviour (no GPU needed) ------------------------------------
import subprocess  # noqa: E402

CLI = os.path.join(ROOT, "pls_b200", "PLS")
needs_cli = pytest.mark.skipif(not os.path.exists(CLI), reason="pls_b200/PLS not built (needs Eigen headers)")


@needs_cli
def test_cli_usage_exits_100():
    r = subprocess.run([CLI], capture_output=True, text=True)                       # main.cpp:12-16
    assert r.returncode == 100 and r.stderr.startswith("Usage: ./pls X_data.csv Y_data.csv num_components")


@needs_cli
def test_cli_ragged_csv_exits_1(tmp_path):
    p = tmp_path / "ragged.csv"
    p.write_text("1,2,3\n4,5\n")
    r = subprocess.run([CLI, str(p), str(p), "1"], capture_output=True, text=True)   # pls.cpp:55-58
    assert r.returncode == 1
    assert "row 1 has 2 columns, but previous row(s) have 3 columns." in r.stderr


@needs_cli
def test_cli_without_gpu_fails_loudly(tmp_path):
    """no CPU fallback: without a usable device the fit exits(1) with the engine's error"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = tmp_path / "a.csv"
    p.write_text("1,2\n3,4\n5,7\n")
    r = subprocess.run([CLI, str(p), str(p), "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "plsb_ctx_create failed" in r.stderr


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the reference's own CPU plsr on a bounded sample) runs without a GPU and prints
    one JSON line with the keys the driver reads."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                          "--ref-rows", "5000"], capture_output=True, text=True, timeout=300, check=True).stdout
    line = json.loads(out.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "pls_fits_per_sec" and line["unit"] == "fits/s"
    assert line["value"] > 0 and line["higher_is_better"] is True and line["dtype"] == "f64"
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] == 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    # both arms print the same `config` dict (the workload only); what differs between them lives in `config_notes`
    assert list(line["config"]) == ["workload"] and line["config_notes"]["sample_rows"] == 5000
