"""Where a small-data CV call (config 3: 1000 LSO folds of 42 x 399, 3 responses) spends its time: PLSB_TRACE laps."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pls_b200 import _capi
from tests.util import load_golden
ctx = _capi.Context(0)
g = load_golden("c3_nir_pls2")
ds = _capi.Dataset(ctx, g["X"], g["Y"])
te = g["lso_test"].astype(np.int64)
for _ in range(3):
    ds.cv_folds(10, 0, te)
os.environ["PLSB_TRACE"] = "1"
for _ in range(2):
    t0 = time.perf_counter(); ds.cv_folds(10, 0, te); print(f"call {1e3*(time.perf_counter()-t0):.3f} ms (traced: includes extra syncs)", flush=True)
del os.environ["PLSB_TRACE"]
E = np.zeros((ds.M, 10, te.size))
for _ in range(3):
    t0 = time.perf_counter(); ds.cv_folds(10, 0, te, out=E); print(f"call with a reused output block {1e3*(time.perf_counter()-t0):.3f} ms", flush=True)
for _ in range(3):
    t0 = time.perf_counter(); ds.cv_press(10, 0, te); print(f"cv_press (residual block stays on the device) {1e3*(time.perf_counter()-t0):.3f} ms", flush=True)
