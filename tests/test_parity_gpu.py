"""GPU parity tests: the CUDA path (through the C ABI) against arrays produced by the unmodified reference.

Tolerances (north_star): one-iteration residual vectors within 1e-12 relative; everything up to the
linear solve is held to 1e-11 or better here, the Krylov solution to 1e-8 (its conditioning amplifies
summation-order differences), CL/CD histories to 1e-6 relative.
"""
import os
import numpy as np
import pytest

from su2_b200 import capi
from tests.parity_util import make_solver, run_first_iteration, run_turb_iteration, load_case, relerr, ROOT

pytestmark = pytest.mark.gpu

CASES = ["naca0012_roe", "naca0012_jst", "box_euler", "rect_ns", "box_tet_roe"]
RANS_CASES = ["box_rans_sa", "rect_rans_sa", "wingbox_rans_sa"]
VARIANTS = ["naca0012_jacobi", "naca0012_fgmres_lusgs", "naca0012_bcgstab_ilu", "naca0012_rfgmres", "naca0012_ls"]
TOL = {  # stage -> relative tolerance
    "default": 1e-12,
    "ilu": 1e-10, "invM": 1e-10,
    "linsyssol": 1e-7, "post_solution": 1e-9,
}


def _tol(tol, key):
    """Tolerance of a report entry; per-variable entries ("stage/var", parity_util.relerr_var) get 1000x the stage's
    (1e-9 where the global measure is held to 1e-12: measured worst case 2.7e-10, the wall-normal momentum residual of the
    stretched wing-box, a component formed by cancellation of terms 1e3 times larger)."""
    if key.endswith("/var"):
        return 1000.0 * tol.get(key[:-4], tol["default"])
    return tol.get(key, tol["default"])


def _gold(name):
    p = os.path.join(ROOT, "tests", "golden", name + ".npz")
    if not os.path.exists(p):
        pytest.skip("no golden fixture " + name)
    return load_case(name)


def _dump_iters(gold):
    its = sorted({int(k[2:k.index("_")]) for k in gold if k.startswith("it") and k.endswith("_pre_solution")})
    return its


@pytest.mark.parametrize("case", CASES)
def test_iteration_stage_by_stage(case):
    """Every stage of one implicit iteration against the reference, at the first and the last dumped iteration."""
    gold = _gold(case)
    its = _dump_iters(gold)
    for it in sorted({its[0], its[-1]}):
        s = make_solver(gold)
        report = {}
        run_first_iteration(s, gold, it, report)
        s.close()
        print(case, "it", it, {k: ("%.2e" % v if isinstance(v, float) else v) for k, v in report.items()})
        bad = {k: v for k, v in report.items() if isinstance(v, float) and not (v <= _tol(TOL, k))}
        assert not bad, (it, bad)
        got, ref = report["linsol_iter"]
        assert got == ref, (it, got, ref)


@pytest.mark.parametrize("case", CASES)
def test_history_tracks_reference(case):
    """rms residuals, CL, CD of the first iterations follow the reference's history."""
    gold = _gold(case)
    s = make_solver(gold)
    n = int(gold["nIter"])
    reps = s.iterate(n)
    s.close()
    hist = gold["history"]
    nv = s.nVar
    for i, r in enumerate(reps):
        ref_rms = hist[i, :nv]
        # components that are identically zero in exact arithmetic (e.g. spanwise momentum) are round-off noise
        ok = np.abs(r["rms_flow"] - ref_rms) <= 1e-6 * np.abs(ref_rms) + 1e-12 * np.max(np.abs(ref_rms))
        assert ok.all(), (i, r["rms_flow"], ref_rms)
        assert abs(r["CL"] - hist[i, 6]) <= 1e-6 * max(1e-3, abs(hist[i, 6])), (i, r["CL"], hist[i, 6])
        assert abs(r["CD"] - hist[i, 7]) <= 1e-6 * max(1e-3, abs(hist[i, 7])), (i, r["CD"], hist[i, 7])
        assert r["linsol_iter_flow"] == int(hist[i, 8]), (i, r["linsol_iter_flow"], hist[i, 8])


@pytest.mark.parametrize("case", RANS_CASES)
def test_rans_sa_stage_by_stage(case):
    """Flow + Spalart-Allmaras iteration against the reference at every dumped iteration: the 3-D hex box, the 2-D
    RECTANGLE (config 3's shape) and the stretched "wing-box" (aspect ratios up to 1e3 at the wall)."""
    gold = _gold(case)
    for it in _dump_iters(gold):
        s = make_solver(gold)
        report = {}
        run_first_iteration(s, gold, it, report)
        run_turb_iteration(s, gold, it, report)
        s.close()
        print(case, "it", it, {k: ("%.2e" % v if isinstance(v, float) else v) for k, v in report.items()})
        tol = dict(TOL, turb_linsyssol=1e-7, turb_post_solution=1e-9, turb_post_mut=1e-8)
        # iteration 0 starts from the uniform free stream: the flow residual is pure round-off (1e-17 of the flux scale)
        # in both codes, so the flow right-hand side and its Krylov solution are not comparable there
        noise = ("bc_residual", "rhs", "res_rms", "linsyssol", "post_solution") if it == 0 else ()
        bad = {k: v for k, v in report.items()
               if isinstance(v, float) and k.split("/")[0] not in noise and not (v <= _tol(tol, k))}
        assert not bad, (it, bad)


@pytest.mark.parametrize("case", RANS_CASES)
def test_rans_sa_history(case):
    gold = _gold(case)
    s = make_solver(gold)
    n = int(gold["nIter"])
    reps = s.iterate(n)
    s.close()
    hist = gold["history"]
    nv = s.nVar
    for i, r in enumerate(reps):
        ref_rms = hist[i, :nv]
        if np.max(np.abs(ref_rms)) < 1e-14:
            continue  # uniform free stream: the flow residual of iteration 0 is round-off in both codes
        ok = np.abs(r["rms_flow"] - ref_rms) <= 1e-6 * np.abs(ref_rms) + 1e-12 * np.max(np.abs(ref_rms))
        assert ok.all(), (i, r["rms_flow"], ref_rms)
        assert abs(r["rms_turb"] - hist[i, 5]) <= 1e-6 * abs(hist[i, 5]), (i, r["rms_turb"], hist[i, 5])
        assert abs(r["CD"] - hist[i, 7]) <= 1e-6 * max(1e-3, abs(hist[i, 7])), (i, r["CD"], hist[i, 7])
        assert r["linsol_iter_flow"] == int(hist[i, 8]) and r["linsol_iter_turb"] == int(hist[i, 10]), (i, r, hist[i])


# ---- Menter SST: two turbulence variables, k in the flow's static energy ---------------------------------
SST_CASES = ["box_rans_sst", "rect_rans_sst", "rect_rans_sst_wls"]


@pytest.mark.parametrize("case", SST_CASES)
def test_rans_sst_stage_by_stage(case):
    """Flow + SST iteration against the reference at every dumped iteration (3-D hex box, 2-D RECTANGLE): flow stages as
    for SA (the primitives now take k out of the static energy), then gradient of (k, omega), convective + diffusive +
    source residual and 2x2 Jacobian after the boundary conditions, Krylov solution, clipped conservative update, and
    F1 / F2 / CDkw / muT of the post-processing."""
    gold = _gold(case)
    for it in _dump_iters(gold):
        s = make_solver(gold)
        report = {}
        run_first_iteration(s, gold, it, report)
        run_turb_iteration(s, gold, it, report)
        s.close()
        print(case, "it", it, {k: ("%.2e" % v if isinstance(v, float) else v) for k, v in report.items()})
        tol = dict(TOL, turb_linsyssol=1e-7, turb_post_solution=1e-9, turb_post_mut=1e-8, sst_F1=1e-9, sst_F2=1e-9, sst_CDkw=1e-8)
        noise = ("bc_residual", "rhs", "res_rms", "linsyssol", "post_solution") if it == 0 else ()
        bad = {k: v for k, v in report.items()
               if isinstance(v, float) and k.split("/")[0] not in noise and not (v <= _tol(tol, k))}
        assert not bad, (it, bad)


@pytest.mark.parametrize("case", SST_CASES)
def test_rans_sst_history(case):
    gold = _gold(case)
    s = make_solver(gold)
    n = int(gold["nIter"])
    reps = s.iterate(n)
    s.close()
    hist = gold["history"]
    nv = s.nVar
    for i, r in enumerate(reps):
        ref_rms = hist[i, :nv]
        if np.max(np.abs(ref_rms)) >= 1e-14:  # iteration 0 of the flow is round-off in both codes
            ok = np.abs(r["rms_flow"] - ref_rms) <= 1e-6 * np.abs(ref_rms) + 1e-12 * np.max(np.abs(ref_rms))
            assert ok.all(), (i, r["rms_flow"], ref_rms)
        assert abs(r["rms_turb"] - hist[i, 5]) <= 1e-6 * abs(hist[i, 5]), (i, r["rms_turb"], hist[i, 5])
        assert abs(r["rms_turb2"] - hist[i, 12]) <= 1e-6 * abs(hist[i, 12]), (i, r["rms_turb2"], hist[i, 12])
        assert abs(r["CD"] - hist[i, 7]) <= 1e-6 * max(1e-3, abs(hist[i, 7])), (i, r["CD"], hist[i, 7])
        assert r["linsol_iter_turb"] == int(hist[i, 10]), (i, r, hist[i])
        if np.max(np.abs(ref_rms)) >= 1e-14:  # a right-hand side of pure round-off: the Krylov count means nothing
            assert r["linsol_iter_flow"] == int(hist[i, 8]), (i, r, hist[i])


# ---- mixed precision: single-precision Jacobian / preconditioner / Krylov vectors --------------------------
@pytest.mark.parametrize("case", ["naca0012_roe_mixed", "box_rans_sa_mixed"])
def test_mixed_precision_follows_the_mixed_reference(case):
    """`mixed_precision = 1` against the reference built with -Denable-mixedprec=true (oracle/build_ref_mixed.sh): the
    nonlinear history. Two single-precision Krylov solves of the same system agree to single-precision round-off amplified
    by the conditioning of the solve (tolerance 1e-4 on the linear residual), not to 1e-12: the gate is the rms history
    within 1 % at every iteration, CL / CD within 1e-4, Krylov iteration counts within one."""
    gold = _gold(case)
    s = make_solver(gold, mixed_precision=1)
    n = int(gold["nIter"])
    reps = s.iterate(n)
    s.close()
    hist = gold["history"]
    nv = s.nVar
    worst = dict(rms=0.0, cl=0.0, cd=0.0, it=0)
    for i, r in enumerate(reps):
        ref = hist[i, :nv]
        if np.max(np.abs(ref)) < 1e-14:
            continue
        worst["rms"] = max(worst["rms"], float(np.max(np.abs(r["rms_flow"] - ref) / np.abs(ref))))
        worst["cl"] = max(worst["cl"], abs(r["CL"] - hist[i, 6]) / max(1e-3, abs(hist[i, 6])))
        worst["cd"] = max(worst["cd"], abs(r["CD"] - hist[i, 7]) / max(1e-3, abs(hist[i, 7])))
        worst["it"] = max(worst["it"], abs(int(r["linsol_iter_flow"]) - int(hist[i, 8])))
    print(case, worst)
    assert worst["rms"] <= 1e-2 and worst["cl"] <= 1e-4 and worst["cd"] <= 1e-4 and worst["it"] <= 1, worst


# ---- code paths that round 1 shipped untested ------------------------------------------------------------------------
@pytest.mark.parametrize("case", VARIANTS)
def test_linear_solver_and_gradient_variants(case):
    """JACOBI, LU_SGS under FGMRES, BCGSTAB+ILU, RESTARTED_FGMRES (restarts every 4) and unweighted LEAST_SQUARES
    reconstruction gradients on the C1 mesh: right-hand side, Krylov solution and iteration count of iteration 0, then
    the rms / CL / CD / iteration-count history, against the reference run with the same options."""
    gold = _gold(case)
    s = make_solver(gold)
    report = {}
    run_first_iteration(s, gold, 0, report)
    s.close()
    print(case, {k: ("%.2e" % v if isinstance(v, float) else v) for k, v in report.items()})
    for k in ("rhs", "linsyssol", "post_solution", "res_rms"):
        assert k in report, (k, "not compared")
    if case == "naca0012_ls":
        assert "grad_recon" in report and "limiter" in report
    tol = dict(TOL)
    bad = {k: v for k, v in report.items() if isinstance(v, float) and not (v <= _tol(tol, k))}
    assert not bad, bad
    got, ref = report["linsol_iter"]
    assert got == ref, (got, ref)
    s = make_solver(gold)
    reps = s.iterate(int(gold["nIter"]))
    s.close()
    _check_history(reps, gold["history"], s.nVar)


def _check_history(reps, hist, nv, rtol=1e-6, iters=True):
    for i, r in enumerate(reps):
        ref_rms = hist[i, :nv]
        ok = np.abs(r["rms_flow"] - ref_rms) <= rtol * np.abs(ref_rms) + 1e-12 * np.max(np.abs(ref_rms))
        assert ok.all(), (i, r["rms_flow"], ref_rms)
        assert abs(r["CL"] - hist[i, 6]) <= rtol * max(1e-3, abs(hist[i, 6])), (i, r["CL"], hist[i, 6])
        assert abs(r["CD"] - hist[i, 7]) <= rtol * max(1e-3, abs(hist[i, 7])), (i, r["CD"], hist[i, 7])
        if iters:
            assert r["linsol_iter_flow"] == int(hist[i, 8]), (i, r["linsol_iter_flow"], hist[i, 8])


@pytest.mark.parametrize("case", ["naca0012_roe", "box_euler", "box_tet_roe", "rect_ns"])
def test_colored_assembly_matches_the_reference(case):
    """ASSEMBLY_COLORED (one launch per edge colour, read-modify-write of residual and diagonal blocks; the reference's
    CFVMFlowSolverBase.inl:1543-1570 strategy) gives the reference's residual and Jacobian like the gather assembly."""
    gold = _gold(case)
    it = _dump_iters(gold)[-1]
    s = make_solver(gold, assembly=capi.ASSEMBLY_COLORED)
    report = {}
    run_first_iteration(s, gold, it, report)
    s.close()
    print(case, "colored", {k: ("%.2e" % v if isinstance(v, float) else v) for k, v in report.items()})
    bad = {k: v for k, v in report.items() if isinstance(v, float) and not (v <= _tol(TOL, k))}
    assert not bad, bad
    assert report["linsol_iter"][0] == report["linsol_iter"][1]


def test_rans_colored_assembly():
    gold = _gold("box_rans_sa")
    s = make_solver(gold, assembly=capi.ASSEMBLY_COLORED)
    reps = s.iterate(int(gold["nIter"]))
    s.close()
    hist = gold["history"]
    for i, r in enumerate(reps[1:], start=1):
        assert np.allclose(r["rms_flow"], hist[i, :5], rtol=1e-6, atol=1e-12 * np.max(np.abs(hist[i, :5]))), (i, r, hist[i])
        assert abs(r["rms_turb"] - hist[i, 5]) <= 1e-6 * abs(hist[i, 5])
        assert r["linsol_iter_flow"] == int(hist[i, 8]) and r["linsol_iter_turb"] == int(hist[i, 10])


def test_prec_partitions_follow_the_hybrid_parallel_reference():
    """prec_partitions = 4 (block-Jacobi ILU0 over four nnz-balanced row partitions) against `SU2_CFD -t 4`, whose
    preconditioner is partitioned the same way (CSysMatrix.cpp:199-227, :701-711): same Krylov iteration counts, same rms
    history.  The reference's 4-thread run sums edge fluxes in another order (reducer strategy), hence 1e-5."""
    gold = _gold("box_rans_sa_t4")
    s = make_solver(gold, prec_partitions=4)
    reps = s.iterate(int(gold["nIter"]))
    s.close()
    hist = gold["history"]
    for i, r in enumerate(reps[1:], start=1):
        assert np.allclose(r["rms_flow"][[0, 1, 2, 4]], hist[i, [0, 1, 2, 4]], rtol=2e-5), (i, r["rms_flow"], hist[i, :5])
        assert abs(r["rms_turb"] - hist[i, 5]) <= 2e-5 * abs(hist[i, 5])
        assert abs(r["CD"] - hist[i, 7]) <= 1e-5 * abs(hist[i, 7])
        assert r["linsol_iter_flow"] == int(hist[i, 8]), (i, r["linsol_iter_flow"], hist[i, 8])
        assert r["linsol_iter_turb"] == int(hist[i, 10]), (i, r["linsol_iter_turb"], hist[i, 10])


# ---- parity gates (ii) and (iii) of SURVEY 8(d): long histories and converged force coefficients -----------------
@pytest.mark.parametrize("case", ["naca0012_roe_long", "naca0012_jst_long", "box_rans_sa_long", "box_rans_sst_long"])
def test_long_history_and_converged_coefficients(case):
    """Gate (ii): the rms history follows the reference's curve (same slope, same order) over the whole run;
    gate (iii): the CONVERGED CL / CD agree to 1e-6 relative.  600 iterations on the C1 mesh take the density residual
    from 1e-1 to below 1e-11 (Roe) with the reference; the RANS box runs 250."""
    gold = _gold(case)
    s = make_solver(gold)
    n = int(gold["nIter"])
    reps = s.iterate(n)
    s.close()
    hist = gold["history"]
    nv = s.nVar
    mine = np.array([r["rms_flow"][0] for r in reps])
    ref = hist[:, 0]
    top = np.max(ref)
    live = ref > 1e-9 * top  # eight orders of reduction (CONV_RESIDUAL_MINVAL -8 of the shipped cfg); below: round-off floor
    dlog = np.abs(np.log10(mine[live]) - np.log10(ref[live]))
    print(case, "max |dlog10 rms[rho]| = %.3e over %d iterations; final rms %.3e vs %.3e; CL %.9f vs %.9f; CD %.9f vs %.9f"
          % (dlog.max(), live.sum(), mine[-1], ref[-1], reps[-1]["CL"], hist[-1, 6], reps[-1]["CD"], hist[-1, 7]))
    assert dlog.max() <= 0.02, dlog.max()          # same curve: within 5 % of the reference at every iteration
    if ref[-1] <= 1e-9 * top:
        assert mine[-1] <= 1e-8 * top              # both runs converged by more than eight orders
    else:
        assert abs(np.log10(mine[-1]) - np.log10(ref[-1])) <= 0.02
    for col, key in ((6, "CL"), (7, "CD")):
        if abs(hist[-1, col]) > 1e-8:
            assert abs(reps[-1][key] - hist[-1, col]) <= 1e-6 * abs(hist[-1, col]), (key, reps[-1][key], hist[-1, col])
    if gold["cfg_rans_sa"] or int(gold.get("cfg_rans_sst", 0)):
        mt = np.array([r["rms_turb"] for r in reps])
        assert np.max(np.abs(np.log10(mt[1:]) - np.log10(hist[1:, 5]))) <= 0.02
    if int(gold.get("cfg_rans_sst", 0)):
        mo = np.array([r["rms_turb2"] for r in reps])
        assert np.max(np.abs(np.log10(mo[1:]) - np.log10(hist[1:, 12]))) <= 0.02


# ---- restart files ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", ["box_euler", "box_rans_sa"])
def test_restart_from_a_reference_file_continues_like_the_reference(case, tmp_path):
    """LoadRestart from the reference's own .dat / .csv file (written after 6 iterations), then 3 iterations: the history
    of the reference restarted from the same file; and a restart written from the device reads back bit for bit."""
    rdir = os.path.join(ROOT, "tests", "golden", "restart")
    gold = _gold(case)
    gid = np.asarray(gold["global_index"]).astype(np.int64)
    ref = np.load(os.path.join(rdir, case + "_restarted_history.npz"))["history"]
    for ext in ("dat", "csv"):
        s = make_solver(gold)
        s.load_restart(os.path.join(rdir, case + "." + ext), gid)
        reps = s.iterate(3)
        nv = s.nVar
        for i, r in enumerate(reps):
            assert np.allclose(r["rms_flow"], ref[i, :nv], rtol=2e-6), (ext, i, r["rms_flow"], ref[i, :nv])
            assert r["linsol_iter_flow"] == int(ref[i, 8]), (ext, i, r["linsol_iter_flow"], ref[i, 8])
            assert abs(r["CD"] - ref[i, 7]) <= 2e-6 * max(1e-3, abs(ref[i, 7]))
            if gold["cfg_rans_sa"]:
                assert abs(r["rms_turb"] - ref[i, 5]) <= 2e-6 * abs(ref[i, 5])
        out = str(tmp_path / ("mine." + ext))
        s.write_restart(out, binary=(ext == "dat"), global_index=gid[:s.nPointDomain])
        sol = s.get("SOLUTION")
        s.close()
        s2 = make_solver(gold)
        s2.load_restart(out, gid)
        back = s2.get("SOLUTION")
        s2.close()
        if ext == "dat":
            assert np.array_equal(back, sol)
        else:
            assert np.allclose(back, sol, rtol=1e-14, atol=0)


def test_cfl_adaptation_follows_the_reference():
    """CFL_ADAPT= YES (CSolver::AdaptCFLNumber): per-point LocalCFL after the first iteration equals the reference's, and
    the 40-iteration history -- the CFL ramps from 10 where the update is not under-relaxed and the linear solver meets its
    tolerance, and is cut where it does not -- follows the reference's (rms, CL/CD 1e-6, equal Krylov counts)."""
    gold = _gold("naca0012_cfl_adapt")
    s = make_solver(gold)
    reps = s.iterate(1)
    cfl = s.get("LOCAL_CFL")
    ref = np.asarray(gold["it0_local_cfl"]).ravel()
    assert np.array_equal(cfl[:s.nPointDomain], ref[:s.nPointDomain]), (np.unique(cfl), np.unique(ref))
    reps += s.iterate(int(gold["nIter"]) - 1)
    final = s.get("LOCAL_CFL")
    s.close()
    print("local CFL after %d iterations: min %.3g max %.3g" % (len(reps), final.min(), final.max()))
    assert final.max() > 10.0
    _check_history(reps, gold["history"], 4)
