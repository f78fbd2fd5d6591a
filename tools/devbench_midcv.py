"""Mid-size cross-validation shapes: too large for the one-launch shared-memory kernel (N*K*8 > ~200 KB), narrower than the
fused fold kernel's range (K < 1024) or with few folds (< 512: one lane, product not split over k).  Prints fits/s and the
engine's kernel times per shape — the numbers DESIGN 3.6 quotes for this regime (ADVICE r2 / VERDICT r2 "mid-size CV").
Usage: python tools/devbench_midcv.py [repeat]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pls_b200 import _capi

ctx = _capi.Context(0)
rep = int(sys.argv[1]) if len(sys.argv) > 1 else 3
SHAPES = [  # N, K, M, A, folds, test fraction (0 = leave-one-out on the first `folds` rows)
    (3000, 1024, 1, 10, 300, 0.0),
    (3000, 1024, 1, 10, 300, 0.3),
    (3000, 1024, 1, 10, 3000, 0.0),
    (2000, 700, 1, 10, 2000, 0.0),       # spectra-sized, cv_LOO
    (2000, 700, 1, 10, 20000, 0.3),      # the CLI's cv_LSO(0.3, 10 N)  (main.cpp:40)
    (2000, 700, 12, 10, 2000, 0.3),      # 8 < M <= 32
]
for N, K, M, A, nf, frac in SHAPES:
    ds = _capi.Dataset(ctx, synth=(20260007, 0, N, K, M, 8))
    if frac > 0:
        _, te = _capi.lso_folds(N, frac, nf, seed=7)
    else:
        te = np.arange(nf, dtype=np.int64).reshape(-1, 1)
    ds.cv_press(A, 0, te)                       # warm-up (builds X'X once; allocations)
    best = 1e30
    ctx.profile(False)                          # wall-clock of the call as a user gets it (profiling changes the lanes)
    for _ in range(rep):
        t0 = time.perf_counter(); ds.cv_press(A, 0, te); best = min(best, time.perf_counter() - t0)
    ctx.profile(3); ctx.profile_reset()         # then one call with an event pair around every launch group
    ds.cv_press(A, 0, te)
    g, ng = ctx.profile_get("cv_gemm"); f, nfo = ctx.profile_get("cv_fold")
    print(f"N={N} K={K} M={M} A={A} folds={nf} test={te.shape[1]}: {best*1e3:.2f} ms = {nf/best:.0f} fits/s | "
          f"product {g:.2f} ms ({ng} launches), fold kernels {f:.2f} ms ({nfo})", flush=True)
    ds.close()
