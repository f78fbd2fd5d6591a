"""Which fused-pass shape is faster on a row shard of a given size UNDER SUSTAINED LOAD (power-capped clocks)?
PLSB_PASS_RB=12: one 512-thread CTA per SM behind a three-tile ring; PLSB_PASS_RB=4: two 256-thread CTAs per SM.
pass.cu picks by size (one-CTA from 6.4 GB up); SCALE_r02 suggested the 4 GB point is on the wrong side on some boards.
Each (size, shape) cell runs for about `secs` seconds, cells interleaved A B A B.  Usage: devbench_shape.py [secs]"""
import os, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pls_b200 import _capi

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 3.0
ctx = _capi.Context(0)
ctx.profile(2)
def clock():
    try:
        return subprocess.check_output(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-i", "0"],
                                       text=True).strip()
    except Exception:
        return "?"
for N in (125_000, 250_000, 500_000, 1_000_000):
    ds = _capi.Dataset(ctx, synth=(20260004, 0, N, 2048, 8, 16))
    for rnd in range(2):
        for rb in ("4", "12"):
            os.environ["PLSB_PASS_RB"] = rb
            ds.fit_resident(32, 0)
            ctx.profile_reset()
            t0 = time.time(); n = 0
            while time.time() - t0 < secs:
                for _ in range(5):
                    ds.fit_resident(32, 0)
                n += 5
            clk = clock()
            tot, nt = ctx.profile_get("fit_total")
            print(f"N={N:8d} ({N*2048*8/1e9:5.1f} GB) shape={'two-CTA' if rb=='4' else 'one-CTA'} round {rnd}: "
                  f"{tot/max(nt,1):.3f} ms per fit over {nt} fits | sm MHz, W at end: {clk}", flush=True)
    os.environ.pop("PLSB_PASS_RB", None)
    ds.close(); ctx.trim()
