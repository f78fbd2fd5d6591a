"""Development benchmark of the batched CV engine: C5-shaped LOO on a subset of folds, C3, C2."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pls_b200 import _capi
from tests.util import load_golden

ctx = _capi.Context(0)
ctx.profile(True)
def report(tag, nf, wall):
    g, ng = ctx.profile_get("cv_gemm"); f, nfo = ctx.profile_get("cv_fold"); x, nx = ctx.profile_get("xtx")
    print(f"{tag}: {nf} folds in {wall*1e3:.1f} ms = {nf/wall:.0f} fits/s | gemm {g:.1f} ms ({ng}) fold-kernels {f:.1f} ms ({nfo}) xtx {x:.1f} ms", flush=True)
    ctx.profile_reset()
for name, A in (("c2_nir_octane", 10), ("c3_nir_pls2", 10)):
    g = load_golden(name)
    ds = _capi.Dataset(ctx, g["X"], g["Y"])
    te = g["lso_test"].astype(np.int64)
    ds.cv_folds(A, 0, te); ctx.profile_reset()
    t0 = time.time(); ds.cv_folds(A, 0, te); report(name + " LSO", te.shape[0], time.time() - t0)
    loo = np.arange(60, dtype=np.int64).reshape(-1, 1)
    t0 = time.time(); ds.cv_folds(A, 0, loo); report(name + " LOO", 60, time.time() - t0)
    ds.close()
N, K, M, A = 20000, 4096, 4, 64
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
ds = _capi.Dataset(ctx, synth=(20260005, 0, N, K, M, 16))
te = np.arange(N, dtype=np.int64).reshape(-1, 1)
ds.cv_folds(A, 0, te, fold0=0, nfolds_local=128); ctx.profile_reset()
t0 = time.time(); ds.cv_folds(A, 0, te, fold0=0, nfolds_local=nf); w = time.time() - t0
report("C5 LOO subset", nf, w)
g_ms = 2.0 * K * K * nf * A / 1e12
print(f"C5: GEMM flops {g_ms:.2f} TFLOP")
