"""dev tool: host-side gaps of the CV engine — wall vs summed kernel time, small calls before/after a big one."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pls_b200 import _capi
from tests.util import load_golden

ctx = _capi.Context(0)
g = load_golden("c2_nir_octane")
loo = np.arange(60, dtype=np.int64).reshape(-1, 1)
ds2 = _capi.Dataset(ctx, g["X"], g["Y"])
def small(tag):
    ds2.cv_folds(10, 0, loo); ctx.sync()
    t0 = time.perf_counter()
    for _ in range(20): ds2.cv_folds(10, 0, loo)
    ctx.sync(); print(f"{tag}: C2 LOO {(time.perf_counter()-t0)/20*1e3:.3f} ms/call", flush=True)
N, K, M, A = 20000, 4096, 4, 64
ds = _capi.Dataset(ctx, synth=(20260005, 0, N, K, M, 16))
te = np.arange(N, dtype=np.int64).reshape(-1, 1)
E = np.zeros((M, A, N))
def big(tag, nf, prof):
    ctx.profile(prof); ctx.profile_reset()
    ctx.sync(); t0 = time.perf_counter()
    ds.cv_folds(A, 0, te, fold0=0, nfolds_local=nf, out=E)
    ctx.sync(); w = time.perf_counter() - t0
    gm, _ = ctx.profile_get("cv_gemm"); f, _ = ctx.profile_get("cv_fold"); x, _ = ctx.profile_get("xtx")
    ctx.profile(False)
    print(f"{tag}: {nf} folds wall {w*1e3:.1f} ms, kernels {gm+f+x:.1f} ms (gemm {gm:.1f} fold {f:.1f}), profile={prof}", flush=True)
small("fresh")
for rep in range(2):
    big("fused", 4096, True); big("fused", 4096, False); big("fused 8192", 8192, True)
    small("after fused")
os.environ["PLSB_CV_V1"] = "1"
for rep in range(2):
    big("v1", 4096, True); big("v1", 4096, False)
    small("after v1")
