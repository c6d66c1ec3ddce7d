"""FAS multigrid: the QuickStart case AS SHIPPED (inv_NACA0012.cfg: JST, FGMRES(10)+ILU0, MGLEVEL= 3 of
which the agglomeration keeps two coarse levels, W-cycle, pre-smoothing 1/2/3) against the unmodified reference.

The agglomerated grids (CMultiGridGeometry) come from the reference's dump like every other geometry in these tests;
everything that runs per iteration -- smoothing on each level with the Lax-Friedrichs scheme on the coarse ones,
restriction, forcing term, prolongation of the correction -- is the library's.  Two tests: every transfer step of the
first W-cycle against the reference's state of each level (tests/golden/naca0012_mg.npz, "it0_ev<n>_<step>_L<level>"), and
the 30-iteration history whose last row is the reference's own regression value (TestCases/serial_regression.py,
euler/naca0012: rms[rho] -3.434541, CL 0.329474, CD 0.021534 -- SURVEY 4.2)."""
import numpy as np
import pytest

from oracle.dumpio import geometry_from_dump, config_from_dump
from su2_b200.solver import FlowSolver, MultiGrid
from tests.parity_util import load_case, relerr

pytestmark = pytest.mark.gpu


def _build(gold):
    nl = int(gold["cfg_mg_levels"])
    levels = []
    for l in range(nl + 1):
        s = FlowSolver(0)
        s.set_geometry(geometry_from_dump(gold, l))
        kw = config_from_dump(gold)
        kw.update(cfl=float(gold["cfg_mg_cfl"][l]), mg_level=l, lax_kappa1=float(gold["cfg_lax_kappa1"]))
        s.configure(**kw)
        s.allocate()
        levels.append(s)
    children = [(gold["mg%d_children_ptr" % l], gold["mg%d_children" % l]) for l in range(1, nl + 1)]
    return MultiGrid(levels, children, cycle=int(gold["cfg_mg_cycle"]), pre_smooth=list(gold["cfg_mg_pre_smooth"]) + [0] * 8,
                     post_smooth=list(gold["cfg_mg_post_smooth"]) + [0] * 8,
                     correction_smooth=list(gold["cfg_mg_correction_smooth"]) + [0] * 8,
                     damp_restriction=float(gold["cfg_mg_damp_restriction"]),
                     damp_prolongation=float(gold["cfg_mg_damp_prolongation"]))


def test_first_w_cycle_step_by_step():
    gold = load_case("naca0012_mg")
    mg = _build(gold)
    nl = len(mg.levels) - 1
    ev = [0]
    worst = {}

    def check(what, k, **arrays):
        name = "it0_ev%d_%s_L%d" % (ev[0], what, k)
        ev[0] += 1
        for key, arr in arrays.items():
            ref = gold[name + "_" + key]  # KeyError = the replay left the reference's sequence of events
            e = relerr(mg.levels[k].get(arr), ref)
            worst[name + "_" + key] = e

    def cycle(k, recursive):
        for _ in range(int(mg.cfg.pre_smooth[k])):
            mg.smooth(k)
        check("presmooth", k, solution="SOLUTION")
        if k < nl:
            f, c = mg.levels[k], mg.levels[k + 1]
            f.Preprocessing(False)
            f.Space_Integration()
            if k > 0:
                mg.SetResidual_Term(k)
            check("fineres", k, linsysres="LINSYSRES")
            mg.SetRestricted_Solution(k)
            c.Preprocessing(False)
            c.Space_Integration()
            check("coarseres", k + 1, solution="SOLUTION", linsysres="LINSYSRES")
            mg.SetForcing_Term(k)
            check("forcing", k + 1, trunc="RES_TRUNC")
            for _ in range(recursive + 1):
                cycle(k + 1, 0 if k == nl - 2 else recursive)
            mg.GetProlongated_Correction(k)
            mg.SmoothProlongated_Correction(k)
            check("correction", k, linsysres="LINSYSRES")
            mg.SetProlongated_Correction(k)
            check("corrected", k, solution="SOLUTION")
            for _ in range(int(mg.cfg.post_smooth[k])):
                mg.smooth(k)

    try:
        mg.levels[0].set("SOLUTION", gold["it0_pre_solution"])
        cycle(0, int(gold["cfg_mg_cycle"]) if int(gold["cfg_mg_cycle"]) < 2 else 0)
        got = mg.levels[0].get("SOLUTION")
    finally:
        mg.close()
    for k, v in worst.items():
        print("  %-44s %.3e" % (k, v))
    n_events = len({k.split("_")[1] for k in gold if k.startswith("it0_ev")})
    assert ev[0] == n_events, (ev[0], n_events)
    # states that went through Krylov solves: 1e-8; the first events (one fine-grid smoothing step) are tighter
    bad = {k: v for k, v in worst.items() if not v <= 1e-8}
    assert not bad, bad
    assert relerr(got, gold["it0_post_solution"]) <= 1e-8


def test_shipped_quickstart_case_follows_the_reference_for_30_iterations():
    gold = load_case("naca0012_mg")
    mg = _build(gold)
    try:
        reps = mg.iterate(int(gold["nIter"]))
    finally:
        mg.close()
    hist = gold["history"]
    for i, r in enumerate(reps):
        ref = hist[i, :4]
        assert np.all(np.abs(r["rms_flow"] - ref) <= 1e-6 * np.abs(ref)), (i, r["rms_flow"], ref)
        assert abs(r["CL"] - hist[i, 6]) <= 1e-6 * abs(hist[i, 6]), (i, r["CL"], hist[i, 6])
        assert abs(r["CD"] - hist[i, 7]) <= 1e-6 * abs(hist[i, 7]), (i, r["CD"], hist[i, 7])
        assert r["linsol_iter_flow"] == int(hist[i, 8])
    last = reps[-1]
    print("iteration 29: log10 rms[rho] %.6f  CL %.6f  CD %.6f" % (np.log10(last["rms_flow"][0]), last["CL"], last["CD"]))
    # the reference's own regression values for this cfg
    assert abs(np.log10(last["rms_flow"][0]) - (-3.434541)) <= 2e-6
    assert abs(last["CL"] - 0.329474) <= 1e-6 and abs(last["CD"] - 0.021534) <= 1e-6
