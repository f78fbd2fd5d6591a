"""Development micro-benchmark (not the driver's bench.py): C4-shaped fit with per-kernel event times
for a few pass-kernel variants.  Usage: python tools/devbench.py [N] [K] [M] [A]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pls_b200 import _capi  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
M = int(sys.argv[3]) if len(sys.argv) > 3 else 8
A = int(sys.argv[4]) if len(sys.argv) > 4 else 32
ctx = _capi.Context(0)
t0 = time.time()
ds = _capi.Dataset(ctx, synth=(20260004, 0, N, K, M, 16))
print(f"synth {N}x{K} in {time.time()-t0:.2f}s", flush=True)
ctx.profile(True)
variants = [dict(), dict(PLSB_STEP_V1="1"), dict(PLSB_PASS_RB="4"), dict(PLSB_PASS_STAGES="3")]
for v in variants:
    for k in ("PLSB_PASS_RB", "PLSB_PASS_STAGES", "PLSB_PASS_CTAS_PER_SM", "PLSB_STEP_V1"):
        os.environ.pop(k, None)
    os.environ.update(v)
    ds.fit_nocopy(A, 0)          # warm
    ctx.profile_reset()
    t0 = time.time()
    reps = 2
    for _ in range(reps):
        ds.fit_nocopy(A, 0)
    ctx.sync()
    wall = (time.time() - t0) / reps
    pf, npf = ctx.profile_get("pass_fused"); px, npx = ctx.profile_get("pass_xty")
    stp, nst = ctx.profile_get("step"); tot, nt = ctx.profile_get("fit_total"); rd, nrd = ctx.profile_get("reduce")
    gb = N * K * 8 / 1e9
    print(f"{v}: fit {wall*1e3:.1f} ms wall, device {tot/max(nt,1):.1f} ms | fused pass {pf/max(npf,1):.3f} ms = "
          f"{gb/(pf/max(npf,1)/1e3):.0f} GB/s | xty {px/max(npx,1):.3f} ms = {gb/(px/max(npx,1)/1e3):.0f} GB/s | "
          f"step {stp/max(nst,1)*1e3:.1f} us | reduce {rd/max(nrd,1)*1e3:.1f} us", flush=True)
if "--type2" in sys.argv:
    ctx.profile_reset()
    ds.fit_nocopy(A, 1)
    ctx.sync()
    x, nx = ctx.profile_get("xtx"); tot, nt = ctx.profile_get("fit_total")
    print(f"type2: fit {tot:.1f} ms, xtx {x:.1f} ms = {2*N*K*K/1e12/(x/1e3):.1f} TFLOP/s (as the reference counts)")
