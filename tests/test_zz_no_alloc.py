"""su2b200_vector_dot_noalloc: the allocation-free variant of su2b200_vector_dot (static device scratch, no cudaMalloc).

It has no counterpart in the reference, so the checker is the existing
entry point su2b200_vector_dot on the same inputs.  Both run the same two reduction kernels with the same grid, hence
the results must agree to the last bit; the gate is the 1e-13 relative tolerance of the other linear-algebra tests.
Two calls in a row with different data: the static array keeps the previous call's contents, the second call must
not see them.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from su2_b200 import capi
from su2_b200.capi import check, ptr_d, as_f64
from su2_b200.solver import LinearSystem

pytestmark = pytest.mark.gpu
TOL = 1e-13


def _chain(P):
    """Tridiagonal pattern on P points (the pattern is irrelevant to the dot product; LinearSystem needs one)."""
    rp = np.zeros(P + 1, dtype=np.int64)
    cols = []
    for i in range(P):
        c = [j for j in (i - 1, i, i + 1) if 0 <= j < P]
        cols.extend(c)
        rp[i + 1] = rp[i] + len(c)
    return rp, np.array(cols, dtype=np.int64)


def _noalloc(ls, a, b):
    a, b = as_f64(a).ravel(), as_f64(b).ravel()
    r = C.c_double(float("nan"))
    rc = ls.lib.su2b200_vector_dot_noalloc(ls.ctx, ptr_d(a), ptr_d(b), C.c_ulong(a.size), C.byref(r))
    return rc, r.value


@pytest.mark.parametrize("nv,P", [(5, 40000), (1, 777), (4, 9)])
def test_noalloc_dot_matches_vector_dot_twice_in_a_row(nv, P):
    rng = np.random.default_rng(20 + nv)
    rp, ci = _chain(P)
    ls = LinearSystem(nv, P, P, rp, ci)
    try:
        n = P * nv
        # second call: shorter effective data (zeros in the tail) and other values, so stale scratch would show
        a1, b1 = rng.standard_normal(n), rng.standard_normal(n)
        a2, b2 = rng.standard_normal(n) * 3.0, rng.standard_normal(n)
        a2[n // 2:] = 0.0
        for a, b in ((a1, b1), (a2, b2)):
            ref = ls.dot(a, b)
            rc, got = _noalloc(ls, a, b)
            assert rc == 0, ls.lib.su2b200_last_error().decode()
            scale = float(np.sum(np.abs(a * b)))
            assert abs(got - ref) <= TOL * scale, (got, ref)
            assert abs(got - float(np.dot(a, b))) <= 1e-12 * scale
        # and once more with the first pair: same answer as before
        rc, again = _noalloc(ls, a1, b1)
        assert rc == 0 and again == _noalloc(ls, a1, b1)[1]
        assert abs(again - ls.dot(a1, b1)) <= TOL * float(np.sum(np.abs(a1 * b1)))
    finally:
        ls.close()


def test_noalloc_dot_limits():
    rp, ci = _chain(4)
    ls = LinearSystem(1, 4, 4, rp, ci)
    try:
        # n == 0 is an empty sum
        z = np.zeros(1)
        r = C.c_double(1.0)
        check(ls.lib.su2b200_vector_dot_noalloc(ls.ctx, ptr_d(z), ptr_d(z), C.c_ulong(0), C.byref(r)))
        assert r.value == 0.0
        # the stated maximum works, one more is rejected with a clear message (no buffer that large is touched)
        nmax = capi.NOALLOC_DOT_MAX
        a = np.full(nmax, 0.5)
        rc, got = _noalloc(ls, a, a)
        assert rc == 0 and abs(got - 0.25 * nmax) <= TOL * 0.25 * nmax
        big = np.zeros(nmax + 1)
        rc, _ = _noalloc(ls, big, big)
        assert rc != 0
        msg = ls.lib.su2b200_last_error().decode()
        assert "SU2B200_NOALLOC_DOT_MAX" in msg and str(nmax) in msg
        # a failed call leaves the entry point usable
        rc, got = _noalloc(ls, a[:100], a[:100])
        assert rc == 0 and got == 25.0
    finally:
        ls.close()
