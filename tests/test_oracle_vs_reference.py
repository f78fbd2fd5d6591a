"""Pin the C restatement directly against the compiled reference (oracle/_ref/libplsref.so) on
random shapes, including ragged / edge cases.  Skipped where the reference was never built."""
import numpy as np
import pytest

from oracle import port, ref
from tests.util import RTOL, assert_state_close, relerr

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (no /root/reference)")


@pytest.mark.parametrize("N,K,M,A", [(40, 7, 1, 5), (33, 12, 2, 6), (200, 31, 5, 9), (17, 17, 3, 4), (64, 5, 8, 5),
                                     (3, 2, 1, 1), (50, 9, 9, 6),
                                     (120, 50, 40, 5), (90, 70, 33, 4)])    # wide response blocks (M > 32)
@pytest.mark.parametrize("method", [0, 1])
def test_random_fit(N, K, M, A, method):
    X, Y = port.synth(1000 + N + K, N, K, M, L=4)
    m = ref.Model(X, Y, A, method)
    assert m.max_imag() == 0.0
    st = port.fit(X, Y, A, method)
    assert_state_close(st, m.state(), rtol=1e-8)
    for c in (1, A):
        assert relerr(port.coefficients(st["R"], st["Q"], c), m.coefficients(c)) <= RTOL * 10
        assert relerr(port.residuals(st["R"], st["Q"], c, X, Y), m.residuals(X, Y, c)) <= RTOL * 10


def test_lso_folds_and_cv():
    X, Y = port.synth(77, 45, 11, 2, L=3)
    tr, te = ref.lso_folds(45, 0.2, 30, seed=42)
    assert sorted(np.concatenate([tr[0], te[0]]).tolist()) == list(range(45))
    E1 = ref.cv_folds(X, Y, 5, 0, te, tr)
    E2 = port.cv_folds(X, Y, 5, 0, te, tr)
    assert relerr(E2, E1) <= RTOL
    m = ref.Model(X, Y, 5, 0)
    assert np.array_equal(m.cv_lso(0.2, 30, seed=42), E1)     # the reference's own driver, same folds
    assert np.array_equal(ref.optimal_num_components(E1), port.optimal_num_components(E2))
    assert relerr(port.validation(E2, 1), ref.validation(E1, 1)) <= RTOL


def test_wilcoxon_random():
    rng = np.random.default_rng(3)
    for n in (5, 18, 60, 301):
        a, b = rng.standard_normal(n), rng.standard_normal(n)
        assert abs(port.wilcoxon(a, b) - ref.wilcoxon(a, b)) < 1e-14
