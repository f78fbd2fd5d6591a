"""Development timings of the less-travelled shapes: wide response blocks (M > 32: step_wide.cuh) and mid-size
cross-validation (data sets between one SM's shared memory and the fused fold kernel's K >= 1024)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pls_b200 import _capi

ctx = _capi.Context(0)
def timed(f, reps=3):
    f(); ctx.sync(); t0 = time.perf_counter()
    for _ in range(reps): f()
    ctx.sync(); return (time.perf_counter() - t0) / reps

for (N, K, M, A) in ((200000, 512, 40, 8), (200000, 512, 8, 8), (200000, 512, 64, 8)):
    ds = _capi.Dataset(ctx, synth=(7, 0, N, K, M, 16))
    ctx.profile(True); ctx.profile_reset()
    t = timed(lambda: ds.fit_resident(A, 0))
    st, ns = ctx.profile_get("step"); pf, npf = ctx.profile_get("pass_fused")
    ctx.profile(False)
    print(f"fit {N}x{K} M={M} A={A}: {t*1e3:.2f} ms per fit | step {st/max(ns,1)*1e3:.1f} us x {ns//4} | pass {pf/max(npf,1)*1e3:.1f} us", flush=True)
    ds.close()

for (N, K, M, A, nf) in ((2000, 500, 2, 10, 2000), (5000, 800, 4, 16, 5000), (3000, 1500, 2, 10, 3000)):
    ds = _capi.Dataset(ctx, synth=(9, 0, N, K, M, 16))
    te = np.arange(N, dtype=np.int64).reshape(-1, 1)[:nf]
    t = timed(lambda: ds.cv_press(A, 0, te), reps=2)
    print(f"LOO {nf} folds of {N}x{K} M={M} A={A}: {t*1e3:.1f} ms = {nf/t:.0f} fits/s", flush=True)
    ds.close()
