"""Pins the numpy restatement of the linear-algebra half of the path (oracle/linalg_oracle.py) against arrays dumped
from the unmodified reference: the Jacobian after the boundary conditions, the time-step diagonal, the ILU0 factors
and inverse diagonal blocks, the right-hand side, the FGMRES solution and its iteration count (tests/golden/*.npz)."""
import os

import numpy as np
import pytest

from oracle import linalg_oracle as lo
from oracle.dumpio import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _system(gold, it=2):
    pre = "it%d_" % it
    rp = np.asarray(gold["row_ptr"]).astype(np.int64)
    ci = np.asarray(gold["col_ind"]).astype(np.int64)
    dp = np.asarray(gold["dia_ptr"]).astype(np.int64)
    A = np.array(gold[pre + "bc_jacobian"], dtype=np.float64)
    nv = A.shape[1]
    vol = np.asarray(gold["volume"], dtype=np.float64)
    dt = np.asarray(gold[pre + "delta_time"], dtype=np.float64)
    Pd = int(gold["nPointDomain"])
    # PrepareImplicitIteration (CFVMFlowSolverBase.hpp:912-973): J_ii += V_i / dt_i
    for i in range(Pd):
        A[dp[i]] += np.eye(nv) * (vol[i] / dt[i])
    return rp, ci, dp, A, nv, Pd


@pytest.mark.parametrize("case", ["box_euler", "box_rans_sa"])
def test_ilu0_restatement_matches_reference_factors(case):
    gold = load_golden(os.path.join(ROOT, "tests", "golden", case + ".npz"))
    rp, ci, dp, A, nv, Pd = _system(gold)
    ilu, invM = lo.ilu0_build(rp, ci, dp, A, Pd)
    ref_ilu = np.asarray(gold["it2_ilu"], dtype=np.float64)
    ref_inv = np.asarray(gold["it2_invM"], dtype=np.float64)
    scale = np.max(np.abs(ref_ilu))
    assert np.max(np.abs(ilu - ref_ilu)) <= 1e-12 * scale
    assert np.max(np.abs(invM - ref_inv)) <= 1e-12 * np.max(np.abs(ref_inv))


@pytest.mark.parametrize("case", ["box_euler", "box_rans_sa"])
def test_fgmres_ilu_restatement_matches_reference_solution(case):
    gold = load_golden(os.path.join(ROOT, "tests", "golden", case + ".npz"))
    rp, ci, dp, A, nv, Pd = _system(gold)
    ilu = np.asarray(gold["it2_ilu"], dtype=np.float64)
    invM = np.asarray(gold["it2_invM"], dtype=np.float64)
    b = np.asarray(gold["it2_sys_linsysres"], dtype=np.float64).ravel()
    P = b.size // nv
    mv = lambda x: lo.block_spmv(rp, ci, A, x, Pd).ravel()
    pc = lambda x: lo.ilu0_apply(rp, ci, dp, ilu, invM, x, Pd).ravel()
    x, it, res = lo.fgmres(mv, pc, b, float(gold["cfg_linear_tol"]), int(gold["cfg_linear_iter"]), Pd * nv)
    ref = np.asarray(gold["it2_linsyssol"], dtype=np.float64).ravel()
    assert it == int(gold["it2_linsol_iter"])
    assert abs(res - float(gold["it2_linsol_res"])) <= 1e-6 * float(gold["it2_linsol_res"])
    assert np.max(np.abs(x - ref)) <= 1e-8 * np.max(np.abs(ref))
    assert x.size == P * nv


def test_matrix_inverse_restatement():
    rng = np.random.default_rng(7)
    for n in (1, 4, 5):
        a = rng.standard_normal((n, n)) + n * np.eye(n)
        assert np.allclose(lo.matrix_inverse(a) @ a, np.eye(n), atol=1e-12)


@pytest.mark.parametrize("variant,prec", [("naca0012_fgmres_lusgs", "lusgs"), ("naca0012_jacobi", "jacobi")])
def test_lusgs_and_jacobi_restatements_match_reference_solution(variant, prec):
    """The LU-SGS / Jacobi restatements inside the pinned FGMRES against the reference's solution of iteration 0 on the
    C1 mesh (same Jacobian as naca0012_roe: only the preconditioner option differs)."""
    base = load_golden(os.path.join(ROOT, "tests", "golden", "naca0012_roe.npz"))
    gold = load_golden(os.path.join(ROOT, "tests", "golden", variant + ".npz"))
    rp, ci, dp, A, nv, Pd = _system(base, it=0)
    b = np.asarray(gold["it0_sys_linsysres"], dtype=np.float64).ravel()
    assert np.array_equal(b, np.asarray(base["it0_sys_linsysres"]).ravel())
    mv = lambda x: lo.block_spmv(rp, ci, A, x, Pd).ravel()
    if prec == "lusgs":
        pc = lambda x: lo.lusgs_apply(rp, ci, dp, A, x, Pd).ravel()
    else:
        inv = np.stack([lo.matrix_inverse(A[dp[i]]) for i in range(Pd)])
        pc = lambda x: np.einsum("ijk,ik->ij", inv, x.reshape(-1, nv)).ravel()
    x, it, res = lo.fgmres(mv, pc, b, float(gold["cfg_linear_tol"]), int(gold["cfg_linear_iter"]), Pd * nv)
    ref = np.asarray(gold["it0_linsyssol"], dtype=np.float64).ravel()
    assert it == int(gold["it0_linsol_iter"])
    assert np.max(np.abs(x - ref)) <= 1e-8 * np.max(np.abs(ref))
