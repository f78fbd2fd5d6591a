"""dev tool: per-phase clock64 offsets inside the step kernel (PLSB_STEP_TRACE=1)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["PLSB_STEP_TRACE"] = "1"
from pls_b200 import _capi
ctx = _capi.Context(0)
ds = _capi.Dataset(ctx, synth=(20260004, 0, 20000, 2048, 8, 16))
ds.fit_nocopy(32, 0)
