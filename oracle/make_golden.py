"""TEST INFRASTRUCTURE ONLY. Generates the golden fixtures under tests/golden/ by running the UNMODIFIED
reference (oracle/_ref/oracle_dump and oracle/_ref/SU2_CFD, built from /root/reference by oracle/build_ref.sh +
build_dump.sh) on the cases under tests/cases/.  Run here (the container with /root/reference); the .npz files are
committed so that the GPU box needs neither the reference nor this script.

usage: python oracle/make_golden.py [case ...]      (no arguments: every case)

Three kinds of fixture:
  full     geometry + config + every intermediate array of the dumped iterations (stage-by-stage parity);
  variant  a base case with a few cfg options replaced (other linear solver / preconditioner / gradient method): config,
           history and the linear solution of iteration 0; the geometry is the base case's (`geometry_from`);
  history  long runs (gates (ii) and (iii) of SURVEY 8(d)): config + history rows only, geometry from the base case;
           `threads` > 1 runs SU2_CFD -t N instead of the dumper (hybrid-parallel reference: block-Jacobi ILU).
Meshes the reference cannot generate itself (tetrahedral box, stretched "wing-box") are written as .su2 files by the
repo's own writer (su2b200_mesh_write_su2) and then read by the reference like any other mesh.
"""
import os
import re
import shutil
import subprocess
import sys
import tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle.dumpio import load_dump  # noqa: E402

REF_MESH_DIR = "/root/reference/QuickStart"

# name -> (cfg, nIter, dump iterations, iteration whose matrices are kept, mesh file or None)
CASES = {
    "naca0012_roe": ("naca0012_roe.cfg", 10, "0", 0, "mesh_NACA0012_inv.su2"),
    "naca0012_jst": ("naca0012_jst.cfg", 10, "0", None, "mesh_NACA0012_inv.su2"),
    "box_euler": ("box_euler.cfg", 6, "0,1,2", 2, None),
    "box_rans_sa": ("box_rans_sa.cfg", 6, "0,1,2", 2, None),
    "rect_ns": ("rect_ns.cfg", 6, "0,1,2", 2, None),
    # round 2: shapes the first five did not cover
    "rect_rans_sa": ("rect_rans_sa.cfg", 6, "0,1,2", 2, None),              # 2-D RANS-SA (config 3's shape)
    "box_tet_roe": ("box_tet_roe.cfg", 6, "0,1,2", 2, "gen:tetbox"),         # 3-D tetrahedra through the flow kernels
    "wingbox_rans_sa": ("wingbox_rans_sa.cfg", 6, "0,1,2", 2, "gen:wingbox"),  # stretched wall-normal spacing
    # Menter SST (V2003m): two-variable turbulence system, k in the flow's static energy
    "box_rans_sst": ("box_rans_sst.cfg", 6, "0,1,2", 2, None),
    "rect_rans_sst": ("rect_rans_sst.cfg", 6, "0,1,2", 2, None),
    "rect_rans_sst_wls": ("rect_rans_sst_wls.cfg", 6, "0,1", None, None),  # weighted least squares for the (k, omega) gradient too
    # FAS multigrid: C1 as shipped (3 coarse levels, W-cycle); agglomerated levels + every transfer step of iteration 0
    "naca0012_mg": ("naca0012_mg.cfg", 30, "0", None, "mesh_NACA0012_inv.su2"),
}
# name -> (base case, {option: value}, nIter)
VARIANTS = {
    "naca0012_jacobi": ("naca0012_roe", {"LINEAR_SOLVER_PREC": "JACOBI"}, 6),
    "naca0012_fgmres_lusgs": ("naca0012_roe", {"LINEAR_SOLVER_PREC": "LU_SGS"}, 6),
    "naca0012_bcgstab_ilu": ("naca0012_roe", {"LINEAR_SOLVER": "BCGSTAB"}, 6),
    "naca0012_rfgmres": ("naca0012_roe", {"LINEAR_SOLVER": "RESTARTED_FGMRES", "LINEAR_SOLVER_ITER": "12",
                                          "LINEAR_SOLVER_RESTART_FREQUENCY": "4"}, 6),
    "naca0012_ls": ("naca0012_roe", {"NUM_METHOD_GRAD_RECON": "LEAST_SQUARES"}, 6),  # unweighted LS is a reconstruction-only option
    # CFL adaptation (CSolver::AdaptCFLNumber) with the shipped QuickStart parameters, from CFL 10; 40 iterations ramp it up
    "naca0012_cfl_adapt": ("naca0012_roe", {"CFL_NUMBER": "10.0", "CFL_ADAPT": "YES",
                                            "CFL_ADAPT_PARAM": "( 0.5, 1.5, 5.0, 400.0 )", "LINEAR_SOLVER_ERROR": "1E-4",
                                            "ITER": "40"}, 40),
}
# name -> (base case, {option: value}, nIter, threads)
HISTORIES = {
    "naca0012_roe_long": ("naca0012_roe", {}, 600, 1),
    "naca0012_jst_long": ("naca0012_jst", {}, 600, 1),
    "box_rans_sa_long": ("box_rans_sa", {}, 250, 1),
    "box_rans_sst_long": ("box_rans_sst", {}, 250, 1),
    "box_rans_sa_t4": ("box_rans_sa", {"MESH_BOX_SIZE": "21,17,13", "LINEAR_SOLVER_ITER": "25"}, 12, 4),
    # the reference built with -Denable-mixedprec=true (oracle/build_ref_mixed.sh): single-precision Jacobian and Krylov solve
    "box_rans_sa_mixed": ("box_rans_sa", {"MESH_BOX_SIZE": "21,17,13"}, 20, 1, "SU2_CFD_mixed"),
    "naca0012_roe_mixed": ("naca0012_roe", {}, 40, 1, "SU2_CFD_mixed"),
}
MATRIX_KEYS = ("_jacobian", "_ilu", "_invM")
KEEP_MATRICES = ("bc_jacobian", "_ilu", "_invM", "turb_bc_jacobian", "turb_src_jacobian", "turb_edge_jacobian", "turb_conv_jacobian")
GEOMETRY_KEYS = ("edge_nodes", "edge_normal", "coord", "volume", "point_ptr", "point_nbr", "point_edge", "row_ptr",
                 "col_ind", "dia_ptr", "edge_ptr", "nNeighbor", "physical_boundary", "wall_distance", "global_index")
VARIANT_KEEP = ("it0_local_cfl", "it0_linsyssol", "it0_post_solution", "it0_linsol_iter", "it0_linsol_res", "it0_sys_linsysres",
                "it0_pre_grad_recon", "it0_pre_limiter", "it0_res_rms", "history", "nIter")


def cfg_with(text, overrides):
    """`KEY= value` lines of `text` with the options in `overrides` replaced (appended when absent)."""
    for k, v in overrides.items():
        pat = re.compile(r"^%s\s*=.*$" % re.escape(k), re.M)
        line = "%s= %s" % (k, v)
        text = pat.sub(line, text) if pat.search(text) else text.rstrip("\n") + "\n" + line + "\n"
    return text


def generate_mesh(kind, path):
    """Meshes the reference's BOX generator cannot make, written by the repo's .su2 writer."""
    from su2_b200.mesh import Mesh
    if kind == "tetbox":
        m = Mesh.box([9, 7, 6], [4.0, 1.0, 1.0], tets=True)
    elif kind == "wingbox":
        m = Mesh.box([11, 13, 7], [2.0, 1.0, 1.0])
        m.stretch(1, 1e-4, 1.6)  # wall-normal (y) spacing: first cell 1e-4 of the height budget, ratio 1.6
    else:
        raise KeyError(kind)
    m.write_su2(path)
    m.close()


def _stage(tmp, cfg_text, cfg_name, mesh):
    with open(os.path.join(tmp, cfg_name), "w") as f:
        f.write(cfg_text)
    if mesh and mesh.startswith("gen:"):
        generate_mesh(mesh[4:], os.path.join(tmp, "mesh.su2"))
    elif mesh:
        shutil.copy(os.path.join(REF_MESH_DIR, mesh), tmp)


def _dump(cfg_text, cfg_name, mesh, nit, dumps):
    tmp = tempfile.mkdtemp(prefix="golden_")
    try:
        _stage(tmp, cfg_text, cfg_name, mesh)
        exe = os.path.join(HERE, "_ref", "oracle_dump")
        subprocess.run([exe, cfg_name, "dump.bin", str(nit), dumps], cwd=tmp, check=True, stderr=subprocess.DEVNULL)
        return load_dump(os.path.join(tmp, "dump.bin"))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def _save(name, d, out_dir=None):
    out_dir = out_dir or os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, name + ".npz")
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in d.items()})
    print(name, "->", path, "%.2f MB" % (os.path.getsize(path) / 1e6))
    return path


def _case_text(name):
    return open(os.path.join(ROOT, "tests", "cases", CASES[name][0])).read()


def run_case(name, keep_all=False, out_dir=None):
    cfg, nit, dumps, keep_it, mesh = CASES[name]
    d = _dump(_case_text(name), cfg, mesh, nit, dumps)
    if not keep_all:
        for k in list(d.keys()):
            if any(k.endswith(m) for m in MATRIX_KEYS):
                keep = keep_it is not None and k.startswith("it%d_" % keep_it) and any(k.endswith(m) for m in KEEP_MATRICES)
                if name.startswith("naca") and not k.endswith("bc_jacobian"):
                    keep = False  # 36k blocks: one copy of the Jacobian is enough for the C1 mesh
                if not keep:
                    del d[k]
    return _save(name, d, out_dir)


def run_variant(name, out_dir=None):
    base, over, nit = VARIANTS[name]
    text = cfg_with(_case_text(base), over)
    with open(os.path.join(ROOT, "tests", "cases", name + ".cfg"), "w") as f:
        f.write("%% variant of %s.cfg generated by oracle/make_golden.py: %s\n" % (base, over) + text)
    d = _dump(text, name + ".cfg", CASES[base][4], nit, "0")
    slim = {k: v for k, v in d.items() if k.startswith(("cfg_", "enum_", "nDim", "nPoint", "nEdge", "nMarker"))
            or k in VARIANT_KEEP}
    slim["geometry_from"] = base
    return _save(name, slim, out_dir)


def _history_from_csv(path):
    """history.csv of SU2_CFD -> rows [rms(5), rms_turb, CL, CD, linsol_iter, linsol_res, turb iters, turb res]."""
    rows = [r for r in open(path).read().splitlines() if r.strip()]
    head = [h.strip().strip('"') for h in rows[0].split(",")]
    col = {h: i for i, h in enumerate(head)}
    out = []
    for r in rows[1:]:
        c = [x.strip() for x in r.split(",")]
        g = lambda key: float(c[col[key]]) if key in col else float("nan")
        rms = [10.0 ** g(k) if k in col else 0.0 for k in ("rms[Rho]", "rms[RhoU]", "rms[RhoV]", "rms[RhoW]", "rms[RhoE]")]
        if "rms[RhoW]" not in col:  # 2-D: energy is the 4th conservative variable
            rms = rms[:3] + [rms[4], 0.0]
        out.append(rms + [10.0 ** g("rms[nu]") if "rms[nu]" in col else 0.0, g("CL"), g("CD"),
                          g("Linear_Solver_Iterations"),  # CFlowCompOutput.cpp:162 names the flow solver's column so
                          10.0 ** g("LinSolRes") if "LinSolRes" in col else float("nan"),
                          g("LinSolIterTurb") if "LinSolIterTurb" in col else 0.0, 0.0])
    return np.array(out)


def run_history(name, out_dir=None):
    base, over, nit, threads, *rest = HISTORIES[name]
    exe = rest[0] if rest else "SU2_CFD"
    text = cfg_with(_case_text(base), dict(over, ITER=str(nit)))
    if threads == 1 and exe == "SU2_CFD":
        d = _dump(text, "case.cfg", CASES[base][4], nit, "")
        slim = {k: v for k, v in d.items() if k.startswith(("cfg_", "enum_", "nDim", "nPoint", "nEdge", "nMarker"))
                or k in ("history", "nIter")}
        if not over:
            slim["geometry_from"] = base
        else:
            slim.update({k: d[k] for k in d if k in GEOMETRY_KEYS or k.startswith("marker")})
        return _save(name, slim, out_dir)
    # hybrid-parallel reference: the real driver with N OpenMP threads; the dumper (1 thread) supplies config values
    text = cfg_with(text, {"HISTORY_OUTPUT": "(ITER, RMS_RES, AERO_COEFF, LINSOL)", "OUTPUT_FILES": "NONE",
                           "CONV_FILENAME": "history", "TABULAR_FORMAT": "CSV", "HISTORY_WRT_FREQ_INNER": "1"})
    d = _dump(text, "case.cfg", CASES[base][4], 1, "")
    tmp = tempfile.mkdtemp(prefix="golden_")
    try:
        _stage(tmp, text, "case.cfg", CASES[base][4])
        env = dict(os.environ, OMP_NUM_THREADS=str(threads))
        subprocess.run([os.path.join(HERE, "_ref", exe), "-t", str(threads), "case.cfg"], cwd=tmp, env=env, check=True,
                       stdout=subprocess.DEVNULL)
        hist = _history_from_csv(os.path.join(tmp, "history.csv"))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    slim = {k: v for k, v in d.items() if k.startswith(("cfg_", "enum_", "nDim", "nPoint", "nEdge", "nMarker", "marker"))
            or k in GEOMETRY_KEYS}
    slim["history"] = hist
    slim["nIter"] = len(hist)
    slim["threads"] = threads
    slim["mixed_precision"] = int(exe == "SU2_CFD_mixed")
    return _save(name, slim, out_dir)


RESTARTS = ("box_euler", "box_rans_sa")


def run_restart(case, out_dir=None):
    """Restart fixtures: the reference's own .dat / .csv restart files after 6 iterations of `case`, and the history of
    the reference RESTARTED from them (3 iterations) -- tests/golden/restart/."""
    out_dir = out_dir or os.path.join(ROOT, "tests", "golden", "restart")
    os.makedirs(out_dir, exist_ok=True)
    text = _case_text(case)
    hist_opts = {"TABULAR_FORMAT": "CSV", "HISTORY_OUTPUT": "(ITER, RMS_RES, AERO_COEFF, LINSOL)", "HISTORY_WRT_FREQ_INNER": "1"}
    a = cfg_with(text, dict(hist_opts, ITER="6", OUTPUT_FILES="(RESTART, RESTART_ASCII)", RESTART_FILENAME="restart_flow",
                            CONV_FILENAME="hist_a"))
    b = cfg_with(text, dict(hist_opts, ITER="3", RESTART_SOL="YES", SOLUTION_FILENAME="restart_flow", OUTPUT_FILES="NONE",
                            CONV_FILENAME="hist_b"))
    tmp = tempfile.mkdtemp(prefix="golden_")
    try:
        exe = os.path.join(HERE, "_ref", "SU2_CFD")
        for name, cfg in (("a.cfg", a), ("b.cfg", b)):
            open(os.path.join(tmp, name), "w").write(cfg)
            subprocess.run([exe, "-t", "1", name], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
        for ext in ("dat", "csv"):
            shutil.copy(os.path.join(tmp, "restart_flow." + ext), os.path.join(out_dir, "%s.%s" % (case, ext)))
        hist = _history_from_csv(os.path.join(tmp, "hist_b.csv"))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    np.savez_compressed(os.path.join(out_dir, case + "_restarted_history.npz"), history=hist)
    print(case, "restart fixtures ->", out_dir)


def run(name, **kw):
    if name.startswith("restart:"):
        return run_restart(name[8:], **kw)
    if name in CASES:
        return run_case(name, **kw)
    if name in VARIANTS:
        return run_variant(name, **kw)
    if name in HISTORIES:
        return run_history(name, **kw)
    raise KeyError(name)


if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES) + list(VARIANTS) + list(HISTORIES) + ["restart:" + c for c in RESTARTS]
    for n in names:
        run(n)
