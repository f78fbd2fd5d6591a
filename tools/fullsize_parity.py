"""Full-size parity against the UNMODIFIED reference (oracle/_ref/libplsref.so, built from /root/reference in the
build container and shipped with the snapshot): the C4 configuration of BASELINE.json — synthetic X 1,000,000 x 2048,
Y x 8, 32 components — fitted on the GPU (KERNEL_TYPE1 streaming passes and KERNEL_TYPE2 DMMA X'X) and by the
reference's own Eigen plsr on the host (single thread, ~2 min, ~35 GB of host memory), same input bits.

    python tools/fullsize_parity.py [--rows 1000000] [--out gpurun_out/fullsize_parity.json]

Dev / evidence tool (the pytest suite checks full size through invariants only because of the reference's run time);
its output is committed under profiles/."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import port, ref                       # noqa: E402  (checker)
from pls_b200 import _capi                         # noqa: E402
from tests.util import col_relerr, relerr, sign_normalise   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "fullsize_parity.json"))
    args = ap.parse_args()
    seed, N, K, M, A = 20260004, args.rows, 2048, 8, 32
    ctx = _capi.Context(0)
    ds = _capi.Dataset(ctx, synth=(seed, 0, N, K, M, 16))
    t0 = time.time(); g1 = ds.fit(A, 0); t1 = time.time() - t0
    g2 = ds.fit(A, 1, want_T=False)
    X, Y = ds.download()
    ds.close()
    t0 = time.time()
    m = ref.Model(X, Y, A, 0)
    want = m.state()
    tref = time.time() - t0
    res = {"config": f"C4 synthetic {N}x{K}, Y x{M}, A={A}, seed {seed}", "reference_fit_seconds_1_thread": tref,
           "gpu_fit_seconds_incl_download_of_results": t1, "max_imag_reference": m.max_imag(), "type1": {}, "type2": {}}
    w = sign_normalise(want)
    Bw = port.coefficients(want["R"], want["Q"], A)
    head = X[:4096]
    Fw = head @ Bw
    for name, got in (("type1", g1), ("type2", g2)):
        g = sign_normalise({k: got.get(k) for k in "WPRQT"})
        for k in "WPRQT":
            if g.get(k) is not None and w.get(k) is not None:
                res[name][k + "_col_relerr"] = col_relerr(g[k], w[k])
        Bg = port.coefficients(got["R"], got["Q"], A)
        res[name]["coefficients_relerr"] = relerr(Bg, Bw)
        res[name]["fitted_head_relerr"] = relerr(head @ Bg, Fw)
        for c in (1, 8, 16, 24):
            res[name][f"coefficients_relerr_c{c}"] = relerr(port.coefficients(got["R"], got["Q"], c),
                                                            port.coefficients(want["R"], want["Q"], c))
    print(json.dumps(res, indent=1))
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)
    ctx.close()


if __name__ == "__main__":
    main()
