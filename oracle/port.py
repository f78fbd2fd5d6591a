"""ctypes binding of oracle/liboracle.so (pls_oracle.c) — TEST INFRASTRUCTURE ONLY.

numpy in / numpy out, column-major (Fortran-order) FP64 like the reference's Eigen matrices.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int64)
i64 = C.c_int64


def build():
    subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.orc_normalcdf.restype = C.c_double
        _lib.orc_normalcdf.argtypes = [C.c_double]
        _lib.orc_wilcoxon.restype = C.c_double
        _lib.orc_synth_u.restype = C.c_double
        _lib.orc_synth_u.argtypes = [C.c_uint64, C.c_uint64]
    return _lib


def F(a):
    return np.asfortranarray(a, dtype=np.float64)


def P(a):
    return a.ctypes.data_as(_dp)


def IP(a):
    return a.ctypes.data_as(_ip)


def fit(X, Y, A, method=0, want_T=True):
    X, Y = F(X), F(Y)
    N, K = X.shape
    M = Y.shape[1]
    W = np.zeros((K, A), order="F")
    Pm = np.zeros((K, A), order="F")
    R = np.zeros((K, A), order="F")
    Q = np.zeros((M, A), order="F")
    T = np.zeros((N, A), order="F") if (want_T and method == 0) else None
    lib().orc_fit(P(X), i64(N), i64(K), P(Y), i64(M), i64(A), C.c_int(method), P(W), P(Pm), P(R), P(Q),
                  P(T) if T is not None else None)
    return dict(W=W, P=Pm, R=R, Q=Q, T=T)


_lib_ext = None


def fit_ext(X, Y, A, want_T=True):
    """The KERNEL_TYPE1 fit in x87 extended precision (oracle/pls_oracle_ext.c): yardstick for FP64 rounding error."""
    global _lib_ext
    if _lib_ext is None:
        so = os.path.join(_HERE, "liboracle_ext.so")
        if not os.path.exists(so):
            raise RuntimeError(f"{so} is not built: run `make -C oracle` (or __graft_entry__.build()) before the tests")
        _lib_ext = C.CDLL(so)
    X, Y = F(X), F(Y)
    N, K = X.shape
    M = Y.shape[1]
    W = np.zeros((K, A), order="F"); Pm = np.zeros((K, A), order="F")
    R = np.zeros((K, A), order="F"); Q = np.zeros((M, A), order="F")
    T = np.zeros((N, A), order="F") if want_T else None
    _lib_ext.orc_fit_ext(P(X), i64(N), i64(K), P(Y), i64(M), i64(A), P(W), P(Pm), P(R), P(Q),
                         P(T) if T is not None else None)
    return dict(W=W, P=Pm, R=R, Q=Q, T=T)


def coefficients(R, Q, comp):
    R, Q = F(R), F(Q)
    K, M = R.shape[0], Q.shape[0]
    B = np.zeros((K, M), order="F")
    lib().orc_coefficients(P(R), P(Q), i64(K), i64(M), i64(comp), P(B))
    return B


def scores(R, comp, Xn):
    R, Xn = F(R), F(Xn)
    S = np.zeros((Xn.shape[0], comp), order="F")
    lib().orc_scores(P(R), i64(R.shape[0]), i64(comp), P(Xn), i64(Xn.shape[0]), P(S))
    return S


def fitted_values(R, Q, comp, Xn):
    R, Q, Xn = F(R), F(Q), F(Xn)
    Fv = np.zeros((Xn.shape[0], Q.shape[0]), order="F")
    lib().orc_fitted_values(P(R), P(Q), i64(R.shape[0]), i64(Q.shape[0]), i64(comp), P(Xn), i64(Xn.shape[0]), P(Fv))
    return Fv


def residuals(R, Q, comp, Xn, Yn):
    R, Q, Xn, Yn = F(R), F(Q), F(Xn), F(Yn)
    E = np.zeros(Yn.shape, order="F")
    lib().orc_residuals(P(R), P(Q), i64(R.shape[0]), i64(Q.shape[0]), i64(comp), P(Xn), i64(Xn.shape[0]), P(Yn), P(E))
    return E


def sse(R, Q, comp, Xn, Yn):
    R, Q, Xn, Yn = F(R), F(Q), F(Xn), F(Yn)
    out = np.zeros(Q.shape[0])
    lib().orc_sse(P(R), P(Q), i64(R.shape[0]), i64(Q.shape[0]), i64(comp), P(Xn), i64(Xn.shape[0]), P(Yn), P(out))
    return out


def sst(X):
    X = F(X)
    out = np.zeros(X.shape[1])
    lib().orc_sst(P(X), i64(X.shape[0]), i64(X.shape[1]), P(out))
    return out


def explained_variance(R, Q, comp, Xn, Yn):
    R, Q, Xn, Yn = F(R), F(Q), F(Xn), F(Yn)
    out = np.zeros(Q.shape[0])
    lib().orc_explained_variance(P(R), P(Q), i64(R.shape[0]), i64(Q.shape[0]), i64(comp), P(Xn), i64(Xn.shape[0]),
                                 P(Yn), P(out))
    return out


def colwise_stdev(X):
    X = F(X)
    out = np.zeros(X.shape[1])
    lib().orc_colwise_stdev(P(X), i64(X.shape[0]), i64(X.shape[1]), P(out))
    return out


def colwise_z_scores(X):
    X = F(X)
    Z = np.zeros(X.shape, order="F")
    lib().orc_colwise_z_scores(P(X), i64(X.shape[0]), i64(X.shape[1]), P(Z))
    return Z


def predict_residuals(R, Q, Xn, Yn):
    """E[y] = (Nn x A) residual matrix of response y for 1..A components (cv_NEW_DATA layout)."""
    R, Q, Xn, Yn = F(R), F(Q), F(Xn), F(Yn)
    K, A = R.shape
    M = Q.shape[0]
    Nn = Xn.shape[0]
    E = np.zeros((M, A, Nn))
    lib().orc_predict_residuals(P(R), P(Q), i64(K), i64(M), i64(A), P(Xn), i64(Nn), P(Yn), P(E))
    return E.transpose(0, 2, 1)  # [M][Nn][A] view of M column-major (Nn x A) blocks


def cv_folds(X, Y, A, method, test_idx, train_idx=None):
    """test_idx: (nfolds, test_size) int64; train_idx: (nfolds, train_size) or None (ascending complement)."""
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
    rows = nfolds * test_size
    E = np.zeros((M, A, rows))
    lib().orc_cv_folds(P(X), i64(N), i64(K), P(Y), i64(M), i64(A), C.c_int(method), i64(nfolds), i64(test_size),
                       IP(test_idx), i64(train_size), IP(train_idx) if train_idx is not None else None, P(E))
    return E.transpose(0, 2, 1)


def _E_in(E):
    """[M][rows][A] -> contiguous M stacked column-major (rows x A) blocks."""
    E = np.asarray(E, dtype=np.float64)
    M, rows, A = E.shape
    return np.ascontiguousarray(E.transpose(0, 2, 1)), M, rows, A


def validation(E, out_type=0):
    buf, M, rows, A = _E_in(E)
    out = np.zeros((M, A), order="F")
    lib().orc_validation(P(buf), i64(M), i64(rows), i64(A), C.c_int(out_type), P(out))
    return out


def normalcdf(z):
    return lib().orc_normalcdf(float(z))


def wilcoxon(e1, e2):
    e1 = np.ascontiguousarray(e1, dtype=np.float64)
    e2 = np.ascontiguousarray(e2, dtype=np.float64)
    return lib().orc_wilcoxon(P(e1), P(e2), i64(e1.size))


def optimal_num_components(E, alpha=0.1):
    buf, M, rows, A = _E_in(E)
    out = np.zeros(M, dtype=np.int64)
    lib().orc_optimal_num_components(P(buf), i64(M), i64(rows), i64(A), C.c_double(alpha), IP(out))
    return out


def synth(seed, N, K, M, L=16, row0=0, rows=None):
    rows = N if rows is None else rows
    X = np.zeros((rows, K), order="F")
    Y = np.zeros((rows, M), order="F")
    lib().orc_synth(C.c_uint64(seed), i64(row0), i64(rows), i64(K), i64(M), i64(L), P(X), i64(rows), P(Y), i64(rows))
    return X, Y


def sign_normalise(state):
    """Flip component columns so the largest-|.| entry of each r_a is positive (W,P,R,Q,T together)."""
    R = state["R"]
    sg = np.sign(R[np.abs(R).argmax(axis=0), np.arange(R.shape[1])])
    sg[sg == 0] = 1.0
    out = {}
    for k, v in state.items():
        out[k] = None if v is None else v * sg[None, :]
    return out
