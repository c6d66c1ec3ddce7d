"""Helpers shared by the parity tests: drive the CUDA path through the C ABI in the reference's call order
and compare every stage with the golden arrays produced by the unmodified reference (oracle/make_golden.py)."""
import os
import numpy as np

from oracle.dumpio import load_golden, geometry_from_dump, config_from_dump
from su2_b200 import capi
from su2_b200.solver import FlowSolver

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def relerr(a, b, floor=1e-12):
    """North-star gate (i): max-norm relative error over the whole array, ||a - b||_inf / ||b||_inf; a reference that is
    pure round-off noise (uniform free stream: residuals of 1e-17) is compared absolutely against the floor."""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.size == 0:
        return 0.0
    scale = max(float(np.max(np.abs(b))), floor)
    return float(np.max(np.abs(a - b)) / scale)


def relerr_var(a, b, floor=1e-12):
    """The same error scaled PER VARIABLE: `b` keeps the reference's shape ([P][nVar], [nnz][nVar][nVar], [P][nG][nDim]),
    every trailing index (a variable, a block entry, a gradient component) is measured against its own largest magnitude
    over the leading axis, so that a density residual cannot hide behind an energy residual five orders larger.
    Trailing indices below 1e-4 of the global scale are round-off of the whole array and use that floor (as does an
    array that is noise altogether: the absolute floor of relerr).  Held to 1000x
    the stage tolerance (a component that is small by cancellation, e.g. the wall-normal momentum residual after the
    symmetry projection, carries the round-off of the large components it was formed from)."""
    b = np.asarray(b, dtype=np.float64)
    a = np.asarray(a, dtype=np.float64)
    if a.size == 0 or b.ndim < 2 or a.size != b.size:
        return relerr(a, b, floor)
    gscale = max(float(np.max(np.abs(b))), floor)
    b2 = b.reshape(b.shape[0], -1)
    a2 = a.reshape(b2.shape)
    scale = np.maximum(np.maximum(np.max(np.abs(b2), axis=0), 1e-4 * gscale), floor)
    return float(np.max(np.max(np.abs(a2 - b2), axis=0) / scale))


def load_case(name):
    """Golden fixture by name; slim fixtures (solver variants, long histories) take the geometry of their base case."""
    gold = load_golden(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    base = gold.get("geometry_from")
    if base:
        bg = load_golden(os.path.join(ROOT, "tests", "golden", str(base) + ".npz"))
        for k, v in bg.items():
            if k not in gold and not k.startswith("it") and k not in ("history", "final_solution"):
                gold[k] = v
        gold["_slim"] = True
    return gold


def make_solver(gold, **overrides):
    s = FlowSolver(0)
    s.set_geometry(geometry_from_dump(gold))
    kw = config_from_dump(gold)
    kw.update(overrides)
    s.configure(**kw)
    s.allocate()
    return s


def init_state_from_gold(s, gold, it):
    """Put the solver in the state the reference had at the start of iteration `it`."""
    pre = "it%d_" % it
    nd, P = s.nDim, s.nPoint
    s.set("SOLUTION", gold[pre + "pre_solution"])
    s.set("SOLUTION_OLD", gold[pre + "pre_solution_old"])
    if s.cfg.rans_sst:
        for name, key in (("SST_F1", "F1"), ("SST_F2", "F2"), ("SST_CDKW", "CDkw")):
            s.set(name, gold[pre + "turb_pre_" + key])
    if s.cfg.rans_sa or s.cfg.rans_sst:
        s.set("EDDY_VISC", gold[pre + "pre_primitive"].reshape(P, nd + 9)[:, nd + 6])
        s.set("TURB_SOLUTION", gold[pre + "turb_pre_solution"])
        s.set("TURB_SOLUTION_OLD", gold[pre + "turb_pre_solution"])


def run_first_iteration(s, gold, it=0, report=None, check=None):
    """One implicit iteration stage by stage; `check(name, got, key)` is called for every comparable array."""
    pre = "it%d_" % it
    nd, P, nv = s.nDim, s.nPoint, s.nVar
    ng = nd + 2
    kinds = [m["kind"] for m in s.markers]

    slim = bool(gold.get("_slim"))

    def cmp(name, got, key, sl=None, optional=False):
        # a missing golden key is an error (a renamed or dropped dump key must not pass silently) unless the fixture is
        # a slim one or the array is one of the matrices make_golden.py keeps for a single iteration only
        if key not in gold:
            if optional or slim:
                return
            raise KeyError("golden fixture lacks %s" % key)
        ref = gold[key]
        if sl is not None:
            got, ref = sl(np.asarray(got)), sl(np.asarray(ref))
        e = relerr(got, ref)
        if report is not None:
            report[name] = e
            report[name + "/var"] = relerr_var(got, ref)
            if "_keep" in report:
                report["_keep"][name] = (np.array(got, copy=True), np.array(ref, copy=True))
        if check is not None:
            check(name, e)

    if it > 0:
        init_state_from_gold(s, gold, it)
    s.Preprocessing(False)
    cmp("primitive", s.get("PRIMITIVE").reshape(P, nd + 9)[:, :nd + 5], pre + "pre_primitive", lambda a: a.reshape(P, -1)[:, :nd + 5])
    if s.cfg.viscous:
        cmp("primitive_visc", s.get("PRIMITIVE").reshape(P, nd + 9), pre + "pre_primitive")
        # nPrimVarGrad = nDim+1 for centred schemes (CEulerVariable.cpp:31-42): the pressure slot is not a gradient there
        nvg = nd + 1 if s.cfg.conv_scheme == capi.CONV_JST else ng
        cmp("grad_prim", s.get("GRAD_PRIM").reshape(P, ng, nd)[:s.nPointDomain, :nvg],
            pre + "pre_grad_prim", lambda a: a.reshape(P, -1, nd)[:s.nPointDomain, :nvg])
    if s.cfg.conv_scheme == capi.CONV_JST:
        cmp("lambda", s.get("LAMBDA"), pre + "pre_lambda")
        cmp("sensor", s.get("SENSOR"), pre + "pre_sensor")
        cmp("und_lapl", s.get("UND_LAPL"), pre + "pre_und_lapl")
    elif s.cfg.muscl:
        cmp("grad_recon", s.get("GRAD_RECON"), pre + "pre_grad_recon")
        if s.cfg.limiter != capi.LIMITER_NONE:
            cmp("limiter", s.get("LIMITER"), pre + "pre_limiter")
    s.Set_OldSolution()
    mn, mx = s.SetTime_Step()
    cmp("delta_time", s.get("DELTA_TIME"), pre + "delta_time")
    cmp("min_dt", [mn], pre + "min_dt")
    cmp("max_dt", [mx], pre + "max_dt")
    s.EdgeFluxResidual()
    cmp("edge_residual", s.get("LINSYSRES"), pre + "edge_linsysres")
    cmp("edge_jacobian", s.get("JACOBIAN"), pre + "edge_jacobian", optional=True)
    for m, k in enumerate(kinds):
        if k == capi.FAR_FIELD:
            s.BC_Far_Field(m)
    cmp("far_residual", s.get("LINSYSRES"), pre + "far_linsysres")
    for m, k in enumerate(kinds):
        if k == capi.HEAT_FLUX:
            s.BC_HeatFlux_Wall(m)
    for m, k in enumerate(kinds):
        if k in (capi.SYMMETRY_PLANE, capi.EULER_WALL):
            s.BC_Sym_Plane(m)
    cmp("bc_residual", s.get("LINSYSRES"), pre + "bc_linsysres")
    cmp("bc_jacobian", s.get("JACOBIAN"), pre + "bc_jacobian", optional=True)
    cmp("bc_solution_old", s.get("SOLUTION_OLD"), pre + "bc_solution_old")
    rms, _, _ = s.PrepareImplicitIteration()
    cmp("rhs", s.get("LINSYSRES"), pre + "sys_linsysres")
    cmp("sys_jacobian", s.get("JACOBIAN"), pre + "sys_jacobian", optional=True)
    cmp("res_rms", rms, pre + "res_rms")
    iters, res = s.Solve()
    if pre + "ilu" in gold:
        cmp("ilu", s.get("ILU"), pre + "ilu")
        cmp("invM", s.get("INVM"), pre + "invM")
    cmp("linsyssol", s.get("LINSYSSOL"), pre + "linsyssol")
    if report is not None:
        report["linsol_iter"] = (iters, int(gold.get(pre + "linsol_iter", -1)))
    s.CompleteImplicitIteration()
    cmp("post_solution", s.get("SOLUTION"), pre + "post_solution")
    cmp("under_relax", s.get("UNDER_RELAX"), pre + "under_relax")
    s.Preprocessing(True)
    return rms


def run_turb_iteration(s, gold, it=0, report=None):
    """CSingleGridIntegration::SingleGrid_Iteration(TURB_SOL) stage by stage, after run_first_iteration(s, gold, it)."""
    pre = "it%d_turb_" % it
    nd, P = s.nDim, s.nPoint

    def cmp(name, got, key, optional=False):
        if key not in gold:
            if optional:
                return
            raise KeyError("golden fixture lacks %s" % key)
        e = relerr(got, gold[key])
        if report is not None:
            report[name] = e
            if "_keep" in report:
                report["_keep"][name] = (np.array(got, copy=True), np.array(gold[key], copy=True))

    s.Turb_Preprocessing()


    nt = 2 if s.cfg.rans_sst else 1
    g = gold[pre + "pre_gradient"].reshape(P, nt * nd)[:s.nPointDomain]
    e = relerr(s.get("TURB_GRADIENT").reshape(P, nt * nd)[:s.nPointDomain], g)
    if report is not None:
        report["turb_gradient"] = e
    s.set("TURB_SOLUTION_OLD", s.get("TURB_SOLUTION"))  # Set_OldSolution
    s.Turb_SetTime_Step()
    cmp("turb_delta_time", s.get("TURB_DELTA_TIME"), pre + "delta_time")
    s.Turb_Space_Integration()
    cmp("turb_bc_residual", s.get("TURB_LINSYSRES"), pre + "bc_linsysres")
    cmp("turb_bc_jacobian", s.get("TURB_JACOBIAN"), pre + "bc_jacobian", optional=True)
    s.Turb_PrepareImplicitIteration()
    iters, res = s.Turb_Solve()
    if report is not None and "history" in gold:
        report["turb_linsol_iter"] = (iters, int(gold["history"][it, 10]))
    cmp("turb_linsyssol", s.get("TURB_LINSYSSOL"), pre + "linsyssol")
    s.Turb_CompleteImplicitIteration()
    s.Turb_Postprocessing()
    cmp("turb_post_solution", s.get("TURB_SOLUTION"), pre + "post_solution")
    cmp("turb_post_mut", s.get("EDDY_VISC"), pre + "post_mut")
    if s.cfg.rans_sst:
        cmp("sst_F1", s.get("SST_F1"), pre + "post_F1")
        cmp("sst_F2", s.get("SST_F2"), pre + "post_F2")
        cmp("sst_CDkw", s.get("SST_CDKW"), pre + "post_CDkw")
    return iters
