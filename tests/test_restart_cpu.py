"""Restart / solution I/O boundary, host side: the SU2 native restart formats read and written by
su2b200_restart_*_arrays, against files written by the unmodified reference (tests/golden/restart/, made by
oracle/make_golden.py restart:<case>) -- and the reference restarted from OUR files where its binary is present."""
import ctypes as C
import os
import re
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

from su2_b200 import capi
from su2_b200.capi import as_f64, as_ulong, check, ptr_d, ptr_ul
from tests.parity_util import load_case, ROOT

RDIR = os.path.join(ROOT, "tests", "golden", "restart")
SU2 = os.path.join(ROOT, "oracle", "_ref", "SU2_CFD")


def _read(path, nd, turb, gid):
    lib = capi.load()
    n = len(gid)
    U = np.zeros(n * (nd + 2))
    nu = np.zeros(n)
    g = as_ulong(gid)
    check(lib.su2b200_restart_read_arrays(path.encode(), nd, int(turb), C.c_ulong(n), ptr_ul(g), ptr_d(U), ptr_d(nu)))
    return U.reshape(n, nd + 2), nu


def _write(path, binary, nd, turb, gid, coord, U, nu):
    lib = capi.load()
    n = len(gid)
    g, c, u, t = as_ulong(gid), as_f64(coord), as_f64(U), as_f64(nu)
    check(lib.su2b200_restart_write_arrays(path.encode(), int(binary), nd, int(turb), C.c_ulong(n), ptr_ul(g), C.c_ulong(n),
                                           ptr_d(c), ptr_d(u), ptr_d(t), 1))


@pytest.mark.parametrize("case", ["box_euler", "box_rans_sa"])
def test_reference_restart_files_are_read_by_global_index(case):
    gold = load_case(case)
    nd, P = int(gold["nDim"]), int(gold["nPoint"])
    turb = bool(gold["cfg_rans_sa"])
    gid = np.asarray(gold["global_index"]).astype(np.int64)  # RCM-permuted local order of the dump
    ref = np.asarray(gold["final_solution"]).reshape(P, nd + 2)
    Ub, nub = _read(os.path.join(RDIR, case + ".dat"), nd, turb, gid)
    Uc, nuc = _read(os.path.join(RDIR, case + ".csv"), nd, turb, gid)
    assert np.array_equal(Ub, ref)                        # the binary file holds the doubles of the run
    assert np.allclose(Uc, ref, rtol=1e-14, atol=0)       # %.15e text
    if turb:
        assert np.allclose(nuc, nub, rtol=1e-14, atol=1e-300) and np.any(nub > 0)


@pytest.mark.parametrize("case", ["box_euler", "box_rans_sa"])
def test_written_restart_is_the_reference_file_byte_for_byte(case, tmp_path):
    gold = load_case(case)
    nd, P = int(gold["nDim"]), int(gold["nPoint"])
    turb = bool(gold["cfg_rans_sa"])
    gid = np.asarray(gold["global_index"]).astype(np.int64)
    coord = np.asarray(gold["coord"]).reshape(P, nd)
    U, nu = _read(os.path.join(RDIR, case + ".dat"), nd, turb, gid)
    out = str(tmp_path / "mine.dat")
    _write(out, True, nd, turb, gid, coord, U, nu)
    assert open(out, "rb").read() == open(os.path.join(RDIR, case + ".dat"), "rb").read()
    out = str(tmp_path / "mine.csv")
    _write(out, False, nd, turb, gid, coord, U, nu)
    mine = [l for l in open(out).read().splitlines() if l.strip()]
    theirs = [l for l in open(os.path.join(RDIR, case + ".csv")).read().splitlines() if l.strip()]
    assert mine[0].replace(" ", "") == theirs[0].replace(" ", "")
    a = np.array([[float(x) for x in l.split(",")] for l in mine[1:]])
    b = np.array([[float(x) for x in l.split(",")] for l in theirs[1:]])
    assert np.array_equal(a, b)
    U2, nu2 = _read(out, nd, turb, gid)
    assert np.allclose(U2, U, rtol=1e-15, atol=0)


@pytest.mark.parametrize("fmt", ["dat", "csv"])
def test_the_reference_restarts_from_our_file(fmt, tmp_path):
    """RESTART_SOL= YES of the unmodified reference on a restart file written HERE continues exactly as from its own."""
    if not os.path.exists(SU2):
        pytest.skip("oracle/_ref/SU2_CFD not built")
    case = "box_euler"
    gold = load_case(case)
    nd, P = int(gold["nDim"]), int(gold["nPoint"])
    gid = np.asarray(gold["global_index"]).astype(np.int64)
    coord = np.asarray(gold["coord"]).reshape(P, nd)
    U, nu = _read(os.path.join(RDIR, case + ".dat"), nd, False, gid)
    tmp = str(tmp_path)
    _write(os.path.join(tmp, "restart_flow." + fmt), fmt == "dat", nd, False, gid, coord, U, nu)
    text = open(os.path.join(ROOT, "tests", "cases", case + ".cfg")).read()
    for k, v in {"ITER": "3", "RESTART_SOL": "YES", "SOLUTION_FILENAME": "restart_flow", "OUTPUT_FILES": "NONE",
                 "CONV_FILENAME": "hist", "TABULAR_FORMAT": "CSV", "HISTORY_OUTPUT": "(ITER, RMS_RES, AERO_COEFF, LINSOL)",
                 "READ_BINARY_RESTART": "YES" if fmt == "dat" else "NO"}.items():
        pat = re.compile(r"^%s\s*=.*$" % re.escape(k), re.M)
        text = pat.sub("%s= %s" % (k, v), text) if pat.search(text) else text + "\n%s= %s\n" % (k, v)
    open(os.path.join(tmp, "case.cfg"), "w").write(text)
    p = subprocess.run([SU2, "-t", "1", "case.cfg"], cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=600)
    assert p.returncode == 0, p.stdout[-2000:]
    rows = [r for r in open(os.path.join(tmp, "hist.csv")).read().splitlines() if r.strip()]
    head = [h.strip().strip('"') for h in rows[0].split(",")]
    got = np.array([[float(x) for x in r.split(",")] for r in rows[1:]])
    ref = np.load(os.path.join(RDIR, case + "_restarted_history.npz"))["history"]
    rms = 10.0 ** got[:, [head.index(k) for k in ("rms[Rho]", "rms[RhoU]", "rms[RhoV]", "rms[RhoW]", "rms[RhoE]")]]
    tol = 1e-8 if fmt == "dat" else 1e-6  # history.csv prints ten digits; the CSV restart carries 16
    assert np.allclose(rms, ref[:len(rms), :5], rtol=tol)
    assert np.array_equal(got[:, head.index("Linear_Solver_Iterations")], ref[:len(rms), 8])
