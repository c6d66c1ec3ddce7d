"""The drop-in boundary, compiled for real: the reference's own SU2_CFD (its driver, CConfig, mesh generator, CGeometry,
output and convergence monitoring) built with the libsu2b200 binding (integration/CB200FlowBackend.hpp, applied to a
scratch copy of the reference by integration/build_shim.sh) and run with ENABLE_CUDA= YES.  The history it writes must be
the history of the unmodified reference (tests/golden/*.npz): rms residuals and CL/CD to 1e-6 relative, equal Krylov
iteration counts.  Geometry and configuration reach the library from the reference's objects, not from mesh_host.cu /
config.py.  Two modes: the whole implicit iteration on the device (default) and only System.Solve (SU2B200_MODE=linear).
The RANS cases (SA and SST) keep the reference's CPU turbulence solver, whose Krylov solver then multiplies through the patched
CSysMatrixVectorProduct::operator() (the reference's own GPU hook) -> su2b200_linsys_spmv."""
import os
import re
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

from tests.parity_util import load_case, ROOT

pytestmark = pytest.mark.gpu
EXE = os.path.join(ROOT, "integration", "_ref", "SU2_CFD_b200")


def _cfg_with(text, overrides):
    for k, v in overrides.items():
        pat = re.compile(r"^%s\s*=.*$" % re.escape(k), re.M)
        line = "%s= %s" % (k, v)
        text = pat.sub(line, text) if pat.search(text) else text.rstrip("\n") + "\n" + line + "\n"
    return text


def _run(case, n_iter, mode):
    text = open(os.path.join(ROOT, "tests", "cases", case + ".cfg")).read()
    text = _cfg_with(text, {"ENABLE_CUDA": "YES", "ITER": str(n_iter), "OUTPUT_FILES": "NONE", "CONV_FILENAME": "history",
                            "TABULAR_FORMAT": "CSV", "HISTORY_OUTPUT": "(ITER, RMS_RES, AERO_COEFF, LINSOL)",
                            "HISTORY_WRT_FREQ_INNER": "1"})
    tmp = tempfile.mkdtemp(prefix="shim_")
    try:
        with open(os.path.join(tmp, "case.cfg"), "w") as f:
            f.write(text)
        env = dict(os.environ, OMP_NUM_THREADS="1")
        if mode == "linear":
            env["SU2B200_MODE"] = "linear"
        else:
            env.pop("SU2B200_MODE", None)
        p = subprocess.run([EXE, "-t", "1", "case.cfg"], cwd=tmp, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                           text=True, timeout=900)
        assert p.returncode == 0, p.stdout[-3000:]
        rows = [r for r in open(os.path.join(tmp, "history.csv")).read().splitlines() if r.strip()]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    head = [h.strip().strip('"') for h in rows[0].split(",")]
    col = {h: i for i, h in enumerate(head)}
    out = []
    for r in rows[1:]:
        c = [x.strip() for x in r.split(",")]
        out.append({h: float(c[i]) for h, i in col.items()})
    return out


@pytest.mark.parametrize("mode", ["iteration", "linear"])
@pytest.mark.parametrize("case", ["box_euler", "rect_ns", "box_rans_sa", "box_rans_sst"])
def test_reference_driver_with_the_binding_reproduces_the_reference_history(case, mode):
    if not os.path.exists(EXE):
        pytest.skip("integration/_ref/SU2_CFD_b200 not built (integration/build_shim.sh, needs /root/reference)")
    gold = load_case(case)
    hist = gold["history"]
    n = int(gold["nIter"])
    rows = _run(case, n, mode)
    assert len(rows) == n
    nd = int(gold["nDim"])
    keys = ["rms[Rho]", "rms[RhoU]", "rms[RhoV]"] + (["rms[RhoW]"] if nd == 3 else []) + ["rms[RhoE]"]
    sst = "cfg_rans_sst" in gold and bool(gold["cfg_rans_sst"])
    for i, r in enumerate(rows):
        ref = hist[i, :nd + 2]
        if sst:  # the reference's own CTurbSSTSolver, fed by the device flow state: (k, omega) residuals and Krylov count
            assert abs(10.0 ** r["rms[k]"] - hist[i, 5]) <= 1e-6 * abs(hist[i, 5]), (i, r["rms[k]"], hist[i, 5])
            assert abs(10.0 ** r["rms[w]"] - hist[i, 12]) <= 1e-6 * abs(hist[i, 12]), (i, r["rms[w]"], hist[i, 12])
            assert int(r["LinSolIterTurb"]) == int(hist[i, 10])
        if np.max(np.abs(ref)) < 1e-14:
            continue  # uniform free stream: round-off residual in both codes
        got = np.array([10.0 ** r[k] for k in keys])
        ok = np.abs(got - ref) <= 1e-6 * np.abs(ref) + 1e-12 * np.max(np.abs(ref))
        assert ok.all(), (i, got, ref)
        assert abs(r["CL"] - hist[i, 6]) <= 1e-6 * max(1e-3, abs(hist[i, 6])), (i, r["CL"], hist[i, 6])
        assert abs(r["CD"] - hist[i, 7]) <= 1e-6 * max(1e-3, abs(hist[i, 7])), (i, r["CD"], hist[i, 7])
        assert int(r["Linear_Solver_Iterations"]) == int(hist[i, 8]), (i, r["Linear_Solver_Iterations"], hist[i, 8])
        if gold["cfg_rans_sa"]:
            assert abs(10.0 ** r["rms[nu]"] - hist[i, 5]) <= 1e-6 * abs(hist[i, 5]), (i, r["rms[nu]"], hist[i, 5])
            assert int(r["LinSolIterTurb"]) == int(hist[i, 10])
