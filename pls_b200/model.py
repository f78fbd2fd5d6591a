"""Python mirror of the reference's PLS::Model interface (include/PLS/pls.h:184-266), on libplsb.so.

Same names, argument meaning and result layouts as the C++ API so the parity tests read like tests of
the reference: Model(X, Y, algorithm, max_components) fits on construction (pls.cpp:340-353); scores /
loadingsX / loadingsY / coefficients / fitted_values / residuals / SSE / explained_variance take an
optional component count; cv_LOO / cv_LSO / cv_NEW_DATA return a Residual whose errors() is a list of
M (rows x A) matrices; validation() and optimal_num_components() are free functions.
All arithmetic happens in the CUDA library (or, for the rank statistics, its host C++).
"""
import numpy as np

from . import _capi
from ._capi import Context, Dataset

KERNEL_TYPE1, KERNEL_TYPE2 = 0, 1       # PLS::METHOD (pls.h:131)
RESS, MSE = 0, 1                        # PLS::VALIDATION_OUTPUT (pls.h:143)

_default_ctx = None


def default_context():
    global _default_ctx
    if _default_ctx is None:
        import os
        _default_ctx = Context(int(os.environ.get("PLSB_DEVICE", "0")))
    return _default_ctx


class Residual:
    """PLS::Residual (pls.h:44-53)."""

    def __init__(self, E, method):
        self._E = np.asarray(E)          # [M][rows][A]
        self._method = method

    def errors(self):
        return [np.asfortranarray(self._E[m]) for m in range(self._E.shape[0])]

    def method(self):
        return self._method

    @property
    def array(self):
        return self._E


def validation(residual, out_type):
    """PLS::validation (pls.cpp:235-261): M x A matrix of RESS or MSE."""
    return _capi.validation(residual.array, out_type)


def optimal_num_components(residual, ALPHA=0.1):
    """PLS::optimal_num_components (pls.cpp:265-289), 1-based, one per response."""
    return _capi.optimal_num_components(residual.array, ALPHA)


class Model:
    def __init__(self, X, Y, algorithm=KERNEL_TYPE1, max_components=None, ctx=None):
        self.ctx = ctx or default_context()
        self._X, self._Y = _capi.F(X), _capi.F(Y)
        if self._Y.ndim == 1:
            self._Y = self._Y.reshape(-1, 1, order="F")
        self.A = int(self._X.shape[1] if max_components is None else max_components)   # pls.cpp:356-359
        self.method = algorithm
        self._ds = None
        self.plsr(self._X, self._Y, algorithm)

    # -- fit ---------------------------------------------------------------------------------------
    def plsr(self, X, Y, algorithm):
        """Model::plsr (pls.cpp:390-437)."""
        if self._ds is not None:
            self._ds.close()
        self._ds = Dataset(self.ctx, X, Y)
        st = self._ds.fit(self.A, algorithm)
        self.W, self.P, self.R, self.Q, self.T = st["W"], st["P"], st["R"], st["Q"], st["T"]
        self.method = algorithm

    # -- queries -----------------------------------------------------------------------------------
    def _comp(self, comp):
        comp = self.A if comp is None else int(comp)
        assert 0 < comp <= self.A
        return comp

    def scores(self, X_new, comp=None):
        return _capi.scores(self.ctx, self.R, self._comp(comp), X_new)

    def loadingsX(self, comp=None):
        return self.P[:, :self._comp(comp)].copy(order="F")

    def loadingsY(self, comp=None):
        return self.Q[:, :self._comp(comp)].copy(order="F")

    def coefficients(self, comp=None):
        comp = self._comp(comp)                      # K x M, tiny: R_c Q_c' (pls.cpp:446)
        return _capi.fitted(self.ctx, self.R, self.Q, comp, np.eye(self.R.shape[0], order="F"))

    def fitted_values(self, X, comp=None):
        return _capi.fitted(self.ctx, self.R, self.Q, self._comp(comp), X)

    def residuals(self, X, Y, comp=None):
        return _capi.fitted(self.ctx, self.R, self.Q, self._comp(comp), X, Y)

    def SSE(self, X, Y, comp=None):
        return _capi.sse(self.ctx, self.R, self.Q, self._comp(comp), X, Y)          # plsb_sse (pls.cpp:457-459)

    def explained_variance(self, X, Y, comp=None):
        Y = _capi.F(Y)
        sst = _capi.colstats(self.ctx, Y)[1] if Y.shape[0] >= 2 else np.zeros(Y.shape[1])   # SST on the device (pls.cpp:69-77)
        return 1.0 - self.SSE(X, Y, comp) / sst

    # -- cross-validation --------------------------------------------------------------------------
    def cv_LOO(self):
        N = self._X.shape[0]
        te = np.arange(N, dtype=np.int64).reshape(-1, 1)
        return Residual(self._ds.cv_folds(self.A, self.method, te), "LOO")

    def cv_LSO(self, test_fraction, num_trials, seed=5489):
        """seed: the std::mt19937 seed (the reference takes the engine itself; default-constructed = 5489)."""
        _, te = _capi.lso_folds(self._X.shape[0], test_fraction, num_trials, seed)
        return Residual(self._ds.cv_folds(self.A, self.method, te), "LSO")

    def cv_NEW_DATA(self, X, Y):
        return Residual(_capi.predict_residuals(self.ctx, self.R, self.Q, X, Y), "NEW DATA")
