"""Development benchmark: LOO folds of the C4 matrix (1M x 2048, Y x 8, A = 32) through the two-lane engine, PRESS on the
device; argv[1] = folds (default 40000).  DEVBENCH_PROF=3 adds event-bracketed kernel sums (inside the overlap),
PLSB_CV_SERIAL=1 the same kernels one after the other."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pls_b200 import _capi

ctx = _capi.Context(0)
N, K, M, A = 1_000_000, 2048, 8, 32
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 40000
ds = _capi.Dataset(ctx, synth=(20260004, 0, N, K, M, 16))
ds.build_gram()
te = np.arange(N, dtype=np.int64).reshape(-1, 1)
ds.cv_press(A, 0, te, fold0=0, nfolds_local=min(nf, 9472))
tag = " ".join(f"{k}={v}" for k, v in sorted(os.environ.items()) if k.startswith("PLSB_"))
if os.environ.get("DEVBENCH_PROF"):
    ctx.profile(int(os.environ["DEVBENCH_PROF"]))
for _ in range(2):
    ctx.profile_reset(); ctx.sync(); t0 = time.perf_counter()
    press = ds.cv_press(A, 0, te, fold0=0, nfolds_local=nf)
    ctx.sync(); w = time.perf_counter() - t0
    print(f"[{tag}] C4 LOO {nf} folds: {w*1e3:.1f} ms = {nf/w:.0f} fits/s (product alone would be {2.0*K*K*A*nf/34.5e12*1e3:.0f} ms at 34.5 TF/s)  press[0,-1] {press[0, -1]:.12e}", flush=True)
    if os.environ.get("DEVBENCH_PROF"):
        g, ng = ctx.profile_get("cv_gemm"); f, nfk = ctx.profile_get("cv_fold")
        print(f"    event-bracketed kernel time: product {g:.1f} ms ({ng} launches), fold kernels {f:.1f} ms ({nfk} launches)", flush=True)
