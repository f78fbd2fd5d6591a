"""Development benchmark of the two-lane CV engine: C5 LOO (all 20000 folds or argv[1]) by the wall clock, no per-kernel
events (profile mode 1 forces one lane).  Variants are chosen with the PLSB_* environment switches of cv.cu / gemm.cu."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pls_b200 import _capi

ctx = _capi.Context(0)
N, K, M, A = 20000, 4096, 4, 64
nf = int(sys.argv[1]) if len(sys.argv) > 1 else N
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ds = _capi.Dataset(ctx, synth=(20260005, 0, N, K, M, 16))
te = np.arange(N, dtype=np.int64).reshape(-1, 1)
E = np.zeros((M, A, N))
ds.cv_folds(A, 0, te, fold0=0, nfolds_local=min(nf, 1024), out=E)
tag = " ".join(f"{k}={v}" for k, v in sorted(os.environ.items()) if k.startswith("PLSB_"))
if os.environ.get("DEVBENCH_PROF"):
    ctx.profile(int(os.environ["DEVBENCH_PROF"]))
for _ in range(reps):
    ctx.profile_reset()
    ctx.sync(); t0 = time.perf_counter()
    ds.cv_folds(A, 0, te, fold0=0, nfolds_local=nf, out=E)
    ctx.sync(); w = time.perf_counter() - t0
    print(f"[{tag}] C5 LOO {nf} folds: {w*1e3:.1f} ms = {nf/w:.0f} fits/s  (GEMM flops {2.0*K*K*nf*A/1e12:.1f} TFLOP -> {2.0*K*K*nf*A/1e12/w:.1f} TF/s if it were all)  checksum {float(np.abs(E[:, :, :nf]).sum()):.12e}", flush=True)
    if os.environ.get("DEVBENCH_PROF"):
        g, ng = ctx.profile_get("cv_gemm"); f, nfk = ctx.profile_get("cv_fold")
        print(f"    event-bracketed kernel time: product {g:.1f} ms ({ng} launches), fold kernels {f:.1f} ms ({nfk} launches)", flush=True)
