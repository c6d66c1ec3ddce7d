"""GPU tests of the CSysMatrix / CSysSolve entry points on sparsity patterns the flow cases do not cover: a
tetrahedral box (irregular rows, fill-in between neighbours -> "unstable" sources and long rows in the ILU kernels)
and a hexahedral box, with random diagonally dominant blocks for nVar = 1, 4, 5.  The checker is the numpy restatement
of the reference routines (oracle/linalg_oracle.py, itself pinned against reference dumps by test_oracle_cpu.py)."""
import numpy as np
import pytest

from oracle import linalg_oracle as lo
from su2_b200 import capi
from su2_b200.mesh import Mesh
from su2_b200.solver import LinearSystem

pytestmark = pytest.mark.gpu


def _pattern(n, tets):
    m = Mesh.box(n, [1.0, 1.0, 1.0], tets=tets)
    for t in ("x_minus", "x_plus", "y_minus", "y_plus", "z_minus", "z_plus"):
        m.set_marker(t, capi.FAR_FIELD)
    m.preprocess()
    rp, ci, dp = (m.array(k).astype(np.int64) for k in ("row_ptr", "col_ind", "dia_ptr"))
    P = m.sizes()["nPoint"]
    m.close()
    return P, rp, ci, dp


def _matrix(rng, rp, ci, dp, nv):
    nnz = int(rp[-1])
    A = rng.standard_normal((nnz, nv, nv)) * 0.3
    for i in range(len(rp) - 1):
        A[dp[i]] += np.eye(nv) * (2.0 + 1.2 * nv * (rp[i + 1] - rp[i]) * 0.3)
    return A


CASES = [([6, 5, 4], True), ([7, 6, 5], False)]


@pytest.mark.parametrize("n,tets", CASES)
@pytest.mark.parametrize("nv", [1, 2, 4, 5])
def test_spmv_ilu_and_solve_match_the_oracle(n, tets, nv):
    rng = np.random.default_rng(100 + nv + (7 if tets else 0))
    P, rp, ci, dp = _pattern(n, tets)
    A = _matrix(rng, rp, ci, dp, nv)
    x = rng.standard_normal(P * nv)
    ls = LinearSystem(nv, P, P, rp, ci)
    try:
        ls.set_matrix(A)
        # MatrixVectorProduct
        y = ls.MatrixVectorProduct(x)
        yr = lo.block_spmv(rp, ci, A, x).ravel()
        assert np.max(np.abs(y - yr)) <= 1e-13 * np.max(np.abs(yr))
        # ILU0 build + apply, every schedule against the oracle and against each other
        ilu, invM = lo.ilu0_build(rp, ci, dp, A)
        zr = lo.ilu0_apply(rp, ci, dp, ilu, invM, x).ravel()
        outs = {}
        for build_mode in (0, 2):
            for apply_mode in (0, 1, 2):
                ls.set_schedule(apply_mode, build_mode)
                ls.BuildPreconditioner(capi.PREC_ILU)
                z = ls.ApplyPreconditioner(capi.PREC_ILU, x)
                outs[(build_mode, apply_mode)] = z
                assert np.max(np.abs(z - zr)) <= 1e-12 * np.max(np.abs(zr)), (build_mode, apply_mode)
        base = outs[(0, 0)]
        for k, z in outs.items():
            assert np.max(np.abs(z - base)) <= 1e-14 * np.max(np.abs(base)), k
        # FGMRES + ILU
        ls.set_schedule(2, 2)
        b = rng.standard_normal(P * nv)
        xs, it, res = ls.Solve(b, capi.SOLVER_FGMRES, capi.PREC_ILU, tol=1e-10, max_iter=30)
        mv = lambda v: lo.block_spmv(rp, ci, A, v).ravel()
        pc = lambda v: lo.ilu0_apply(rp, ci, dp, ilu, invM, v).ravel()
        xo, ito, reso = lo.fgmres(mv, pc, b, 1e-10, 30)
        assert it == ito
        assert np.max(np.abs(xs - xo)) <= 1e-9 * np.max(np.abs(xo))
        assert np.max(np.abs(mv(xs) - b)) <= 1e-8 * np.max(np.abs(b))
    finally:
        ls.close()


def test_in_place_preconditioner_application_and_repeated_builds():
    """x == y at the C ABI (the polling sweeps need distinct buffers internally) and rebuilding with a new matrix."""
    rng = np.random.default_rng(5)
    P, rp, ci, dp = _pattern([6, 5, 4], True)
    ls = LinearSystem(5, P, P, rp, ci)
    try:
        for rep in range(2):
            A = _matrix(rng, rp, ci, dp, 5)
            ls.set_matrix(A)
            ls.BuildPreconditioner(capi.PREC_ILU)
            x = rng.standard_normal(P * 5)
            ilu, invM = lo.ilu0_build(rp, ci, dp, A)
            zr = lo.ilu0_apply(rp, ci, dp, ilu, invM, x).ravel()
            z = ls.ApplyPreconditioner(capi.PREC_ILU, x)
            assert np.max(np.abs(z - zr)) <= 1e-12 * np.max(np.abs(zr))
    finally:
        ls.close()


@pytest.mark.parametrize("n,tets", CASES)
@pytest.mark.parametrize("nv", [2, 4, 5])
@pytest.mark.parametrize("wide,mode", [(1, 1), (1, 2), (8, 1), (40, 1)])
def test_wide_level_kernels_match_the_oracle(n, tets, nv, wide, mode):
    """The kernels that carry the benchmark -- k_poll_sweep_rows (wide-level ILU0 application) and k_factor_rows (wide-level
    factorisation) -- only take levels of at least 4096 rows by default, wider than any level of a test mesh.  Here the
    threshold is lowered (su2b200_linsys_set_wide_threshold) so that they run on the small tetrahedral and hexahedral
    boxes: threshold 1 = every level, 8 / 40 = hybrid with a narrow head and tail; factor_mode 2 = rows kernel only."""
    rng = np.random.default_rng(300 + nv + wide + (7 if tets else 0))
    P, rp, ci, dp = _pattern(n, tets)
    A = _matrix(rng, rp, ci, dp, nv)
    x = rng.standard_normal(P * nv)
    ilu, invM = lo.ilu0_build(rp, ci, dp, A)
    zr = lo.ilu0_apply(rp, ci, dp, ilu, invM, x).ravel()
    ls = LinearSystem(nv, P, P, rp, ci)
    try:
        ls.set_matrix(A)
        ls.set_schedule(2, 2)
        ls.BuildPreconditioner(capi.PREC_ILU)
        z_default = ls.ApplyPreconditioner(capi.PREC_ILU, x)
        ls.set_wide_threshold(wide, wide, mode)
        ls.BuildPreconditioner(capi.PREC_ILU)
        z = ls.ApplyPreconditioner(capi.PREC_ILU, x)
        assert np.max(np.abs(z - zr)) <= 1e-12 * np.max(np.abs(zr))
        assert np.max(np.abs(z - z_default)) <= 1e-14 * np.max(np.abs(zr))
        b = rng.standard_normal(P * nv)
        xs, it, res = ls.Solve(b, capi.SOLVER_FGMRES, capi.PREC_ILU, tol=1e-10, max_iter=30)
        mv = lambda v: lo.block_spmv(rp, ci, A, v).ravel()
        pc = lambda v: lo.ilu0_apply(rp, ci, dp, ilu, invM, v).ravel()
        xo, ito, reso = lo.fgmres(mv, pc, b, 1e-10, 30)
        assert it == ito
        assert np.max(np.abs(xs - xo)) <= 1e-9 * np.max(np.abs(xo))
    finally:
        ls.close()


def test_stale_slim_factors_are_refused():
    """BuildILUPreconditioner / ComputeILUPreconditioner are separate entry points: with slim factors (U left in the
    Jacobian) an application after the matrix changed would mix the old L, D with the new U -- it must fail loudly."""
    rng = np.random.default_rng(11)
    P, rp, ci, dp = _pattern([7, 6, 5], False)
    ls = LinearSystem(5, P, P, rp, ci)
    try:
        ls.set_matrix(_matrix(rng, rp, ci, dp, 5))
        ls.BuildPreconditioner(capi.PREC_ILU)
        x = rng.standard_normal(P * 5)
        ls.ApplyPreconditioner(capi.PREC_ILU, x)
        ls.set_matrix(_matrix(rng, rp, ci, dp, 5))
        with pytest.raises(capi.Su2B200Error, match="stale"):
            ls.ApplyPreconditioner(capi.PREC_ILU, x)
        ls.BuildPreconditioner(capi.PREC_ILU)
        ls.ApplyPreconditioner(capi.PREC_ILU, x)
    finally:
        ls.close()


def test_jacobi_and_lusgs_against_the_oracle():
    """Jacobi (CSysMatrix.cpp:648-670) and LU-SGS (:826-888) preconditioners on the tetrahedral pattern."""
    rng = np.random.default_rng(21)
    P, rp, ci, dp = _pattern([6, 5, 4], True)
    for nv in (4, 5):
        A = _matrix(rng, rp, ci, dp, nv)
        x = rng.standard_normal(P * nv)
        ls = LinearSystem(nv, P, P, rp, ci)
        try:
            ls.set_matrix(A)
            ls.BuildPreconditioner(capi.PREC_JACOBI)
            z = ls.ApplyPreconditioner(capi.PREC_JACOBI, x).reshape(P, nv)
            zr = np.stack([np.linalg.solve(A[dp[i]], x.reshape(P, nv)[i]) for i in range(P)])
            assert np.max(np.abs(z - zr)) <= 1e-12 * np.max(np.abs(zr))
            z = ls.ApplyPreconditioner(capi.PREC_LU_SGS, x).reshape(P, nv)
            zr = lo.lusgs_apply(rp, ci, dp, A, x).reshape(P, nv)
            assert np.max(np.abs(z - zr)) <= 1e-12 * np.max(np.abs(zr))
        finally:
            ls.close()


@pytest.mark.parametrize("n,tets", CASES)
@pytest.mark.parametrize("nv", [2, 4, 5])
def test_precondition_and_multiply_equals_the_two_operators(n, tets, nv):
    """z = M^-1 v, w = A z in one call (the fused upper product on the hexahedral pattern, the plain sequence on the
    tetrahedral one, where fill-in modifies upper blocks) against ApplyPreconditioner + MatrixVectorProduct and the
    oracle. The fused product adds the upper part of a row as one pre-summed term: one rounding per row apart."""
    rng = np.random.default_rng(300 + nv + (7 if tets else 0))
    P, rp, ci, dp = _pattern(n, tets)
    A = _matrix(rng, rp, ci, dp, nv)
    ls = LinearSystem(nv, P, P, rp, ci)
    try:
        ls.set_matrix(A)
        ilu, invM = lo.ilu0_build(rp, ci, dp, A)
        for prec in (capi.PREC_ILU, capi.PREC_JACOBI):
            ls.BuildPreconditioner(prec)
            for rep in range(2):  # twice: the second call finds the buffers of the first
                v = rng.standard_normal(P * nv)
                z, w = ls.PreconditionAndMultiply(prec, v)
                z2 = ls.ApplyPreconditioner(prec, v)
                w2 = ls.MatrixVectorProduct(z2)
                assert np.array_equal(z, z2)
                assert np.max(np.abs(w - w2)) <= 1e-14 * np.max(np.abs(w2))
                if prec == capi.PREC_ILU:
                    zr = lo.ilu0_apply(rp, ci, dp, ilu, invM, v).ravel()
                    wr = lo.block_spmv(rp, ci, A, zr).ravel()
                    assert np.max(np.abs(z - zr)) <= 1e-13 * np.max(np.abs(zr))
                    assert np.max(np.abs(w - wr)) <= 1e-13 * np.max(np.abs(wr))
    finally:
        ls.close()
