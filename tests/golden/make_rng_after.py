"""tests/golden/lso_rng_after.json: the state cv_LSO leaves the caller's generator in (pls.cpp:512, 534), taken from the
UNMODIFIED reference (oracle/_ref/libplsref.so) for the three (data set, trials) pairs tests/test_gpu_cpp_api.py runs.
A default-constructed std::mt19937 (seed 5489) goes in, as in tests/cpp/api_check.cpp; the value stored is its next
output.  Run here (needs /root/reference): python tests/golden/make_rng_after.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import ref
from tests.util import load_golden

out = {}
for name, A, trials in [("c1_toy", 2, 100), ("c2_nir_octane", 10, 600), ("c3_nir_pls2", 10, 1000)]:
    g = load_golden(name)
    m = ref.Model(g["X"], g["Y"], A, 0)
    out[f"{name}:{trials}"] = m.cv_lso_rng_after(0.3, trials)
    m.close()
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "lso_rng_after.json"), "w"), indent=1)
print(out)
