"""pls_b200 — B200-native (sm_100a) engine for the PLS fit / cross-validation path of tjhladish/PLS.

The product is libplsb.so (hand-written CUDA behind the C-ABI of include/plsb.h) and the C++ drop-in
layer (include/PLS/pls.h + pls_b200/host/).  This Python package is the ctypes binding used by the
tests and bench.py, plus `Model`, a mirror of PLS::Model's interface for scripting.
"""
from . import _capi
from ._capi import Context, Dataset, PlsbError, device_count, fit_host, lib
from .model import KERNEL_TYPE1, KERNEL_TYPE2, MSE, RESS, Model, Residual, optimal_num_components, validation

__all__ = ["Context", "Dataset", "PlsbError", "device_count", "fit_host", "lib", "Model", "Residual", "validation",
           "optimal_num_components", "KERNEL_TYPE1", "KERNEL_TYPE2", "RESS", "MSE", "_capi"]
