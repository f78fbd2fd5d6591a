"""ctypes binding of oracle/_ref/libplsref.so — the UNMODIFIED reference (tjhladish/PLS) compiled
from /root/reference by oracle/Makefile.  TEST INFRASTRUCTURE ONLY.

`available()` is False where neither /root/reference nor a prebuilt oracle/_ref/ exists.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from .port import F, P, IP, i64, _E_in

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libplsref.so")
CLI = os.path.join(_HERE, "_ref", "PLS_ref")
_lib = None


def build():
    if os.path.exists("/root/reference/src/pls.cpp"):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)


def available():
    if not os.path.exists(_SO):
        try:
            build()
        except Exception:
            return False
    return os.path.exists(_SO)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libplsref.so not built (needs /root/reference)")
        _lib = C.CDLL(_SO)
        _lib.plsref_model_new.restype = C.c_void_p
        _lib.plsref_model_max_imag.restype = C.c_double
        _lib.plsref_wilcoxon.restype = C.c_double
        _lib.plsref_normalcdf.restype = C.c_double
        _lib.plsref_normalcdf.argtypes = [C.c_double]
        _lib.plsref_time_fit.restype = C.c_double
        _lib.plsref_read_matrix_file.restype = C.c_int64
    return _lib


class Model:
    """PLS::Model(X, Y, method, A) of the reference (pls.h:187-191)."""

    def __init__(self, X, Y, A, method=0, omp=False):
        """omp=True drives the -fopenmp build of the same sources (Eigen threads its GEMM kernels only)."""
        self.X, self.Y = F(X), F(Y)
        self.N, self.K = self.X.shape
        self.M = self.Y.shape[1]
        self.A, self.method = A, method
        self._lib = lib_omp() if omp else lib()
        self.h = C.c_void_p(self._lib.plsref_model_new(P(self.X), i64(self.N), i64(self.K), P(self.Y), i64(self.M),
                                                       C.c_int(method), i64(A)))

    def close(self):
        if getattr(self, "h", None):
            self._lib.plsref_model_free(self.h)
            self.h = None

    __del__ = close

    def state(self):
        K, M, A, N = self.K, self.M, self.A, self.N
        W = np.zeros((K, A), order="F"); Pm = np.zeros((K, A), order="F")
        R = np.zeros((K, A), order="F"); Q = np.zeros((M, A), order="F")
        T = np.zeros((N, A), order="F") if self.method == 0 else None
        self._lib.plsref_model_state(self.h, P(W), P(Pm), P(R), P(Q), P(T) if T is not None else None)
        return dict(W=W, P=Pm, R=R, Q=Q, T=T)

    def max_imag(self):
        return self._lib.plsref_model_max_imag(self.h)

    def scores(self, Xn, comp):
        Xn = F(Xn)
        S = np.zeros((Xn.shape[0], comp), order="F")
        self._lib.plsref_scores(self.h, P(Xn), i64(Xn.shape[0]), i64(self.K), i64(comp), P(S))
        return S

    def coefficients(self, comp):
        B = np.zeros((self.K, self.M), order="F")
        self._lib.plsref_coefficients(self.h, i64(comp), P(B))
        return B

    def fitted_values(self, Xn, comp):
        Xn = F(Xn)
        Fv = np.zeros((Xn.shape[0], self.M), order="F")
        self._lib.plsref_fitted_values(self.h, P(Xn), i64(Xn.shape[0]), i64(self.K), i64(self.M), i64(comp), P(Fv))
        return Fv

    def residuals(self, Xn, Yn, comp):
        Xn, Yn = F(Xn), F(Yn)
        E = np.zeros(Yn.shape, order="F")
        self._lib.plsref_residuals(self.h, P(Xn), i64(Xn.shape[0]), i64(self.K), P(Yn), i64(self.M), i64(comp), P(E))
        return E

    def sse(self, Xn, Yn, comp):
        Xn, Yn = F(Xn), F(Yn)
        out = np.zeros(self.M)
        self._lib.plsref_sse(self.h, P(Xn), i64(Xn.shape[0]), i64(self.K), P(Yn), i64(self.M), i64(comp), P(out))
        return out

    def explained_variance(self, Xn, Yn, comp):
        Xn, Yn = F(Xn), F(Yn)
        out = np.zeros(self.M)
        self._lib.plsref_explained_variance(self.h, P(Xn), i64(Xn.shape[0]), i64(self.K), P(Yn), i64(self.M), i64(comp),
                                        P(out))
        return out

    def cv_loo(self):
        """The reference's own cv_LOO (inner models extract K components: slow)."""
        E = np.zeros((self.M, self.A, self.N))
        self._lib.plsref_cv_loo(self.h, P(E))
        return E.transpose(0, 2, 1)

    def cv_lso(self, frac, trials, seed=5489):
        test_size = int(frac * self.N + 0.5)
        E = np.zeros((self.M, self.A, trials * test_size))
        self._lib.plsref_cv_lso(self.h, C.c_double(frac), i64(trials), C.c_uint32(seed), P(E))
        return E.transpose(0, 2, 1)

    def cv_lso_rng_after(self, frac, trials, seed=5489):
        """the next output of the caller's std::mt19937 after the reference's cv_LSO has drawn its folds from it"""
        self._lib.plsref_cv_lso_rng_after.restype = C.c_uint32
        return int(self._lib.plsref_cv_lso_rng_after(self.h, C.c_double(frac), i64(trials), C.c_uint32(seed)))

    def cv_new_data(self, Xn, Yn):
        Xn, Yn = F(Xn), F(Yn)
        E = np.zeros((self.M, self.A, Xn.shape[0]))
        self._lib.plsref_cv_new_data(self.h, P(Xn), i64(Xn.shape[0]), i64(self.K), P(Yn), i64(self.M), P(E))
        return E.transpose(0, 2, 1)


def lso_folds(N, frac, trials, seed=5489):
    test_size = int(frac * N + 0.5)
    train = np.zeros((trials, N - test_size), dtype=np.int64)
    test = np.zeros((trials, test_size), dtype=np.int64)
    lib().plsref_lso_folds(i64(N), C.c_double(frac), i64(trials), C.c_uint32(seed), IP(train), IP(test))
    return train, test


def cv_folds(X, Y, A, method, test_idx, train_idx=None, out=None, row0=0):
    """Per-fold refit with the reference's plsr at inner A (bit-identical to the first A of inner K)."""
    X, Y = F(X), F(Y)
    N, K = X.shape
    M = Y.shape[1]
    test_idx = np.ascontiguousarray(test_idx, dtype=np.int64)
    nfolds, test_size = test_idx.shape
    if train_idx is not None:
        train_idx = np.ascontiguousarray(train_idx, dtype=np.int64)
        train_size = train_idx.shape[1]
    else:
        train_size = N - test_size
    if out is None:
        out = np.zeros((M, A, nfolds * test_size))
    rows = out.shape[2]
    lib().plsref_cv_folds(P(X), i64(N), i64(K), P(Y), i64(M), C.c_int(method), i64(A), i64(nfolds), i64(test_size),
                          IP(test_idx), i64(train_size), IP(train_idx) if train_idx is not None else None, P(out),
                          i64(rows), i64(row0))
    return out.transpose(0, 2, 1)


def validation(E, out_type=0):
    buf, M, rows, A = _E_in(E)
    out = np.zeros((M, A), order="F")
    lib().plsref_validation(P(buf), i64(M), i64(rows), i64(A), C.c_int(out_type), P(out))
    return out


def optimal_num_components(E, alpha=0.1):
    buf, M, rows, A = _E_in(E)
    out = np.zeros(M, dtype=np.int64)
    lib().plsref_optimal_num_components(P(buf), i64(M), i64(rows), i64(A), C.c_double(alpha), IP(out))
    return out


def wilcoxon(e1, e2):
    e1 = np.ascontiguousarray(e1, dtype=np.float64)
    e2 = np.ascontiguousarray(e2, dtype=np.float64)
    return lib().plsref_wilcoxon(P(e1), P(e2), i64(e1.size))


def normalcdf(z):
    return lib().plsref_normalcdf(float(z))


def colwise_z_scores(X):
    X = F(X)
    Z = np.zeros(X.shape, order="F")
    lib().plsref_colwise_z_scores(P(X), i64(X.shape[0]), i64(X.shape[1]), P(Z))
    return Z


def colwise_stdev(X):
    X = F(X)
    out = np.zeros(X.shape[1])
    lib().plsref_colwise_stdev(P(X), i64(X.shape[0]), i64(X.shape[1]), P(out))
    return out


def sst(X):
    X = F(X)
    out = np.zeros(X.shape[1])
    lib().plsref_sst(P(X), i64(X.shape[0]), i64(X.shape[1]), P(out))
    return out


def read_matrix_file(path):
    r, c = C.c_int64(), C.c_int64()
    n = lib().plsref_read_matrix_file(path.encode(), C.byref(r), C.byref(c), None, i64(0))
    out = np.zeros((r.value, c.value), order="F")
    lib().plsref_read_matrix_file(path.encode(), C.byref(r), C.byref(c), P(out), i64(n))
    return out


def time_fit(X, Y, A, method=0, reps=1):
    X, Y = F(X), F(Y)
    return lib().plsref_time_fit(P(X), i64(X.shape[0]), i64(X.shape[1]), P(Y), i64(Y.shape[1]), C.c_int(method),
                                 i64(A), C.c_int(reps))


_SO_OMP = os.path.join(_HERE, "_ref", "libplsref_omp.so")
_lib_omp = None


def _bind(l):
    l.plsref_model_new.restype = C.c_void_p
    l.plsref_model_max_imag.restype = C.c_double
    l.plsref_wilcoxon.restype = C.c_double
    l.plsref_normalcdf.restype = C.c_double
    l.plsref_normalcdf.argtypes = [C.c_double]
    l.plsref_time_fit.restype = C.c_double
    l.plsref_read_matrix_file.restype = C.c_int64
    return l


def lib_omp():
    """The -fopenmp build (None-safe callers check omp_available())."""
    global _lib_omp
    if _lib_omp is None:
        if not os.path.exists(_SO_OMP):
            raise RuntimeError("oracle/_ref/libplsref_omp.so not built")
        _lib_omp = _bind(C.CDLL(_SO_OMP))
    return _lib_omp


def omp_available():
    return os.path.exists(_SO_OMP)


def time_fit_omp(X, Y, A, method=0, reps=1):
    """time_fit on the -fopenmp build of the same reference sources (Eigen threads its GEMM kernels only);
    None when that build is absent."""
    if not omp_available():
        return None
    X, Y = F(X), F(Y)
    return lib_omp().plsref_time_fit(P(X), i64(X.shape[0]), i64(X.shape[1]), P(Y), i64(Y.shape[1]), C.c_int(method),
                                    i64(A), C.c_int(reps))
