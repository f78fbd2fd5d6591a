"""dev tool: time the DMMA scores kernel (S = X R) against the FMA kernel on a C4-shaped slice."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pls_b200 import _capi

ctx = _capi.Context(0)
rng = np.random.default_rng(1)
N, K = int(os.environ.get("ROWS", 200000)), 2048
X = np.asfortranarray(rng.standard_normal((N, K)))
for A in (32, 8, 13):
    R = np.asfortranarray(rng.standard_normal((K, A)))
    want = X @ R
    for fma in ("0", "1"):
        os.environ["PLSB_SCORES_FMA"] = fma
        ctx.profile(True); ctx.profile_reset()
        S = _capi.scores(ctx, R, A, X)
        ms, n = ctx.profile_get("scores")
        err = np.abs(S - want).max() / np.abs(want).max()
        print(f"A={A} fma={fma} relerr={err:.2e} dmma_kernel_ms={ms:.3f} launches={n} "
              f"GB/s={(8.0 * N * K) / (ms * 1e-3) / 1e9 if ms else 0:.0f} TF/s={(2.0 * N * K * A) / (ms * 1e-3) / 1e12 if ms else 0:.1f}")
