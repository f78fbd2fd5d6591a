"""ctypes binding of pls_b200/libplsb.so (include/plsb.h).  No torch, no numpy arithmetic: numpy arrays
are only the host containers handed to the C-ABI, exactly as Eigen matrices are in the C++ layer.

There is no CPU fallback: if the library is missing, or no sm_100 device is visible, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PLSB_LIB") or os.path.join(_HERE, "libplsb.so")   # PLSB_LIB: dev A/B builds of the same library
_lib = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int64)
i64 = C.c_int64

# symbols include/plsb.h declares (tests check that the library exports every one)
SYMBOLS = [
    "plsb_last_error", "plsb_version", "plsb_device_count", "plsb_ctx_create", "plsb_ctx_destroy", "plsb_ctx_sync", "plsb_ctx_trim",
    "plsb_comm_unique_id", "plsb_comm_init", "plsb_comm_allreduce_test", "plsb_comm_info",
    "plsb_dataset_create", "plsb_dataset_create_device", "plsb_dataset_create_synth", "plsb_dataset_destroy",
    "plsb_dataset_shape", "plsb_dataset_concat", "plsb_dataset_download", "plsb_dataset_set_sharded", "plsb_dataset_build_gram",
    "plsb_fit", "plsb_fit_resident", "plsb_model_download", "plsb_model_download_scores", "plsb_model_destroy", "plsb_fit_host", "plsb_xty", "plsb_xtx",
    "plsb_scores", "plsb_fitted", "plsb_predict_residuals", "plsb_coefficients", "plsb_sse_prefixes", "plsb_sse",
    "plsb_cv_folds", "plsb_cv_folds_press", "plsb_colstats", "plsb_colstats_static", "plsb_colwise_z_scores", "plsb_dataset_zscore",
    "plsb_normalcdf", "plsb_wilcoxon", "plsb_validation", "plsb_optimal_num_components", "plsb_lso_folds",
    "plsb_profile_enable", "plsb_profile_reset", "plsb_profile_get", "plsb_launch_count",
]

PROF = dict(pass_fused=0, pass_xty=1, step=2, xtx=3, cv_gemm=4, cv_fold=5, fit_total=6, upload=7, reduce=8, scores=9)


class PlsbError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PlsbError(f"{LIB_PATH} is not built (run `make` or __graft_entry__.build()); there is no CPU fallback")
        _lib = C.CDLL(LIB_PATH)
        _lib.plsb_last_error.restype = C.c_char_p
        _lib.plsb_version.restype = C.c_char_p
        _lib.plsb_normalcdf.restype = C.c_double
        _lib.plsb_normalcdf.argtypes = [C.c_double]
        _lib.plsb_wilcoxon.restype = C.c_double
    return _lib


def check(rc):
    if rc != 0:
        raise PlsbError(f"plsb error {rc}: {lib().plsb_last_error().decode()}")


def F(a):
    return np.asfortranarray(a, dtype=np.float64)


def P(a):
    return None if a is None else a.ctypes.data_as(_dp)


def IP(a):
    return a.ctypes.data_as(_ip)


def device_count():
    n = C.c_int(0)
    check(lib().plsb_device_count(C.byref(n)))
    return n.value


class Context:
    def __init__(self, device=0):
        self.h = C.c_void_p()
        check(lib().plsb_ctx_create(C.c_int(device), C.byref(self.h)))
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            lib().plsb_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def sync(self):
        check(lib().plsb_ctx_sync(self.h))

    def trim(self):
        check(lib().plsb_ctx_trim(self.h))

    # -- communicator (row-sharded fits) -------------------------------------------------------
    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        check(lib().plsb_comm_unique_id(buf))
        return bytes(buf.raw)

    def comm_init(self, rank, nranks, uid):
        check(lib().plsb_comm_init(self.h, C.c_int(rank), C.c_int(nranks), C.c_char_p(uid)))

    def comm_info(self):
        r, n, p = C.c_int(0), C.c_int(1), C.c_int(0)
        check(lib().plsb_comm_info(self.h, C.byref(r), C.byref(n), C.byref(p)))
        return {"rank": r.value, "nranks": n.value, "exchange": "peer" if p.value else ("nccl" if n.value > 1 else "none")}

    def allreduce_test(self, arr):
        arr = np.ascontiguousarray(arr, dtype=np.float64)
        check(lib().plsb_comm_allreduce_test(self.h, P(arr), i64(arr.size)))
        return arr

    # -- profiling -----------------------------------------------------------------------------
    def profile(self, on=True):
        """on: False/0 off, True/1 events around every launch group, 2 only around whole fits, 3 like 1 with the CV
        engine's two lanes kept (kernels timed inside the overlap)"""
        check(lib().plsb_profile_enable(self.h, C.c_int(int(on) if on in (2, 3) else (1 if on else 0))))

    def profile_reset(self):
        check(lib().plsb_profile_reset(self.h))

    def profile_get(self, slot):
        ms, n = C.c_double(0), C.c_int64(0)
        check(lib().plsb_profile_get(self.h, C.c_int(PROF[slot] if isinstance(slot, str) else slot), C.byref(ms),
                                     C.byref(n)))
        return ms.value, n.value

    def launch_count(self):
        n = C.c_int64(0)
        check(lib().plsb_launch_count(self.h, C.byref(n)))
        return n.value


class Dataset:
    """Device-resident (X, Y): the role of Model's _X,_Y copies (pls.cpp:344)."""

    def __init__(self, ctx, X=None, Y=None, *, synth=None, sharded=False):
        self.ctx = ctx
        self.h = C.c_void_p()
        if synth is not None:
            seed, row0, N, K, M, L = synth
            check(lib().plsb_dataset_create_synth(ctx.h, C.c_uint64(seed), i64(row0), i64(N), i64(K), i64(M), i64(L),
                                                  C.byref(self.h)))
            self.N, self.K, self.M = N, K, M
        else:
            X, Y = F(X), F(Y)
            if Y.ndim == 1:
                Y = Y.reshape(-1, 1, order="F")
            self.N, self.K = X.shape
            self.M = Y.shape[1]
            assert Y.shape[0] == self.N
            check(lib().plsb_dataset_create(ctx.h, P(X), i64(max(self.N, 1)), i64(self.N), i64(self.K), P(Y),
                                            i64(max(self.N, 1)), i64(self.M), C.byref(self.h)))
        if sharded:
            check(lib().plsb_dataset_set_sharded(self.h, C.c_int(1)))

    @classmethod
    def concat(cls, ctx, parts):
        """the rows of `parts` (data sets on any GPUs) one after the other, as a new data set on ctx's device"""
        self = cls.__new__(cls)
        self.ctx, self.h = ctx, C.c_void_p()
        arr = (C.c_void_p * len(parts))(*[p.h for p in parts])
        check(lib().plsb_dataset_concat(ctx.h, i64(len(parts)), arr, C.byref(self.h)))
        self.N, self.K, self.M = sum(p.N for p in parts), parts[0].K, parts[0].M
        return self

    def close(self):
        if getattr(self, "h", None):
            lib().plsb_dataset_destroy(self.h)
            self.h = None

    __del__ = close

    def download(self):
        X = np.zeros((self.N, self.K), order="F")
        Y = np.zeros((self.N, self.M), order="F")
        check(lib().plsb_dataset_download(self.h, P(X), i64(max(self.N, 1)), P(Y), i64(max(self.N, 1))))
        return X, Y

    def zscore(self):
        xm = np.zeros(self.K); xs = np.zeros(self.K); ym = np.zeros(self.M); ys = np.zeros(self.M)
        check(lib().plsb_dataset_zscore(self.h, P(xm), P(xs), P(ym), P(ys)))
        return xm, xs, ym, ys

    def fit(self, A, method=0, want_T=True):
        K, M, N = self.K, self.M, self.N
        W = np.zeros((K, A), order="F"); Pm = np.zeros((K, A), order="F")
        R = np.zeros((K, A), order="F"); Q = np.zeros((M, A), order="F")
        T = np.zeros((N, A), order="F") if (want_T and method == 0) else None
        check(lib().plsb_fit(self.h, i64(A), C.c_int(method), P(W), P(Pm), P(R), P(Q), P(T)))
        return dict(W=W, P=Pm, R=R, Q=Q, T=T)

    def fit_nocopy(self, A, method=0):
        """fit without reading results back and WITHOUT forming the scores T (A/B only: the reference's KERNEL_TYPE1
        fit stores T, so throughput figures use fit_resident)"""
        check(lib().plsb_fit(self.h, i64(A), C.c_int(method), None, None, None, None, None))

    def fit_resident(self, A, method=0, download=None):
        """the complete fit (W, P, R, Q and, KERNEL_TYPE1, the N x A scores T) formed and kept in HBM;
        download: None, or a string of matrix names ("WPRQ", "T", ...) to read back before the model is released"""
        mh = C.c_void_p()
        check(lib().plsb_fit_resident(self.h, i64(A), C.c_int(method), C.byref(mh)))
        out = None
        try:
            if download:
                K, M, N = self.K, self.M, self.N
                out = {k: np.zeros((K, A), order="F") for k in "WPR" if k in download}
                if "Q" in download:
                    out["Q"] = np.zeros((M, A), order="F")
                if "T" in download:
                    out["T"] = np.zeros((N, A), order="F")
                check(lib().plsb_model_download(mh, *[P(out.get(k)) for k in "WPRQT"]))
                if "C" in download:   # the scores as the drop-in's Model::T takes them: (re, im) pairs, im = 0
                    Tc = np.zeros((N, A), order="F", dtype=np.complex128)
                    check(lib().plsb_model_download_scores(mh, Tc.ctypes.data_as(_dp), C.c_int(1)))
                    out["C"] = Tc
        finally:
            lib().plsb_model_destroy(mh)
        return out

    def xty(self):
        S = np.zeros((self.K, self.M), order="F")
        check(lib().plsb_xty(self.h, P(S)))
        return S

    def xtx(self, download=True):
        G = np.zeros((self.K, self.K), order="F") if download else None
        check(lib().plsb_xtx(self.h, P(G)))
        return G

    def build_gram(self, across_ranks=False):
        """X'X and X'Y built once and kept on the data set (collective when across_ranks: every rank reads its row slice)."""
        check(lib().plsb_dataset_build_gram(self.h, C.c_int(1 if across_ranks else 0)))

    def cv_folds(self, A, method, test_idx, fold0=0, nfolds_local=None, out=None):
        """test_idx: (nfolds_total, test_size) int64 rows left out per fold; folds [fold0, fold0+nfolds_local)
        are computed by this call (fold sharding).  Returns E[M][nfolds_total*test_size][A]."""
        test_idx = np.ascontiguousarray(test_idx, dtype=np.int64)
        ntot, ts = test_idx.shape
        nloc = ntot - fold0 if nfolds_local is None else nfolds_local
        if out is None:
            out = np.zeros((self.M, A, ntot * ts))
        check(lib().plsb_cv_folds(self.h, i64(A), C.c_int(method), i64(ntot), i64(ts), IP(test_idx), i64(fold0),
                                  i64(nloc), P(out)))
        return out.transpose(0, 2, 1)

    def cv_press(self, A, method, test_idx, fold0=0, nfolds_local=None, want_E=False):
        """PRESS sums (M x A) of this call's folds reduced on the device; the residual block stays there unless want_E."""
        test_idx = np.ascontiguousarray(test_idx, dtype=np.int64)
        ntot, ts = test_idx.shape
        nloc = ntot - fold0 if nfolds_local is None else nfolds_local
        press = np.zeros((self.M, A), order="F")
        E = np.zeros((self.M, A, ntot * ts)) if want_E else None
        check(lib().plsb_cv_folds_press(self.h, i64(A), C.c_int(method), i64(ntot), i64(ts), IP(test_idx), i64(fold0),
                                        i64(nloc), P(E), P(press)))
        return (press, E.transpose(0, 2, 1)) if want_E else press


def colstats(ctx, X, means=None):
    X = F(X); N, K = X.shape
    mean = np.zeros(K); sst = np.zeros(K); sd = np.zeros(K)
    mi = None if means is None else np.ascontiguousarray(means, dtype=np.float64)
    check(lib().plsb_colstats(ctx.h, P(X), i64(max(N, 1)), i64(N), i64(K), P(mi), P(mean), P(sst), P(sd)))
    return mean, sst, sd


def colstats_static(device, stream, dX, ldx, N, K, d_mean=None, d_sst=None, d_stdev=None):
    """plsb_colstats_static: the allocation-free variant.  Every pointer is a raw DEVICE address (int) that the caller
    owns — the column-major N x K matrix and up to three K-vectors of results; `stream` is a cudaStream_t handle (0 =
    default stream).  Nothing is copied or synchronised here: the results are ready in stream order."""
    vp = lambda a: C.c_void_p(a) if a else None
    check(lib().plsb_colstats_static(C.c_int(device), vp(stream), vp(dX), i64(ldx), i64(N), i64(K), vp(d_mean), vp(d_sst),
                                     vp(d_stdev)))


def colwise_z_scores(ctx, X, mean=None, stdev=None):
    X = F(X); N, K = X.shape
    Z = np.zeros((N, K), order="F")
    m = None if mean is None else np.ascontiguousarray(mean, dtype=np.float64)
    s = None if stdev is None else np.ascontiguousarray(stdev, dtype=np.float64)
    check(lib().plsb_colwise_z_scores(ctx.h, P(X), i64(max(N, 1)), i64(N), i64(K), P(m), P(s), P(Z), i64(max(N, 1))))
    return Z


def fit_host(ctx, X, Y, A, method=0, want_T=True):
    X, Y = F(X), F(Y)
    N, K = X.shape
    M = Y.shape[1]
    W = np.zeros((K, A), order="F"); Pm = np.zeros((K, A), order="F")
    R = np.zeros((K, A), order="F"); Q = np.zeros((M, A), order="F")
    T = np.zeros((N, A), order="F") if (want_T and method == 0) else None
    check(lib().plsb_fit_host(ctx.h, P(X), i64(max(N, 1)), i64(N), i64(K), P(Y), i64(max(N, 1)), i64(M), i64(A),
                              C.c_int(method), P(W), P(Pm), P(R), P(Q), P(T)))
    return dict(W=W, P=Pm, R=R, Q=Q, T=T)


def scores(ctx, R, comp, Xn):
    R, Xn = F(R), F(Xn)
    S = np.zeros((Xn.shape[0], comp), order="F")
    check(lib().plsb_scores(ctx.h, P(R), i64(R.shape[0]), i64(comp), P(Xn), i64(max(Xn.shape[0], 1)), i64(Xn.shape[0]),
                            P(S)))
    return S


def fitted(ctx, R, Q, comp, Xn, Yn=None):
    R, Q, Xn = F(R), F(Q), F(Xn)
    Nn = Xn.shape[0]
    Yn = None if Yn is None else F(Yn)
    out = np.zeros((Nn, Q.shape[0]), order="F")
    check(lib().plsb_fitted(ctx.h, P(R), P(Q), i64(R.shape[0]), i64(Q.shape[0]), i64(comp), P(Xn), i64(max(Nn, 1)),
                            i64(Nn), P(Yn), i64(max(Nn, 1)), P(out)))
    return out


def sse(ctx, R, Q, comp, Xn, Yn):
    """Model::SSE (pls.cpp:457-459): column sums of squared residuals with `comp` components, reduced on the device"""
    R, Q, Xn, Yn = F(R), F(Q), F(Xn), F(Yn)
    Nn = Xn.shape[0]
    out = np.zeros(Q.shape[0])
    check(lib().plsb_sse(ctx.h, P(R), P(Q), i64(R.shape[0]), i64(Q.shape[0]), i64(comp), P(Xn), i64(max(Nn, 1)), i64(Nn),
                         P(Yn), i64(max(Nn, 1)), P(out)))
    return out


def predict_residuals(ctx, R, Q, Xn, Yn):
    R, Q, Xn, Yn = F(R), F(Q), F(Xn), F(Yn)
    K, A = R.shape
    M = Q.shape[0]
    Nn = Xn.shape[0]
    E = np.zeros((M, A, Nn))
    check(lib().plsb_predict_residuals(ctx.h, P(R), P(Q), i64(K), i64(M), i64(A), P(Xn), i64(max(Nn, 1)), i64(Nn),
                                       P(Yn), i64(max(Nn, 1)), P(E)))
    return E.transpose(0, 2, 1)


# ---- host-side statistics (C++ in libplsb.so; no GPU involved) ------------------------------------
def _E_in(E):
    """[M][rows][A] -> contiguous M stacked column-major (rows x A) blocks."""
    E = np.asarray(E, dtype=np.float64)
    M, rows, A = E.shape
    return np.ascontiguousarray(E.transpose(0, 2, 1)), M, rows, A


def normalcdf(z):
    return lib().plsb_normalcdf(float(z))


def wilcoxon(e1, e2):
    e1 = np.ascontiguousarray(e1, dtype=np.float64)
    e2 = np.ascontiguousarray(e2, dtype=np.float64)
    return lib().plsb_wilcoxon(P(e1), P(e2), i64(e1.size))


def validation(E, out_type=0):
    buf, M, rows, A = _E_in(E)
    out = np.zeros((M, A), order="F")
    check(lib().plsb_validation(P(buf), i64(M), i64(rows), i64(A), C.c_int(out_type), P(out)))
    return out


def optimal_num_components(E, alpha=0.1):
    buf, M, rows, A = _E_in(E)
    out = np.zeros(M, dtype=np.int64)
    check(lib().plsb_optimal_num_components(P(buf), i64(M), i64(rows), i64(A), C.c_double(alpha), IP(out)))
    return out


def lso_folds(N, frac, trials, seed=5489):
    test_size = int(frac * N + 0.5)
    test = np.zeros((trials, test_size), dtype=np.int64)
    train = np.zeros((trials, N - test_size), dtype=np.int64)
    check(lib().plsb_lso_folds(i64(N), C.c_double(frac), i64(trials), C.c_uint32(seed), IP(test), IP(train)))
    return train, test
