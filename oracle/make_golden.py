"""TEST INFRASTRUCTURE ONLY. Generates the golden fixtures under tests/golden/ by running the UNMODIFIED
reference (oracle/_ref/oracle_dump, built from /root/reference by oracle/build_ref.sh + build_dump.sh)
on the cases under tests/cases/.  Run here (the container with /root/reference); the .npz files are
committed so that the GPU box needs neither the reference nor this script.

usage: python oracle/make_golden.py [case ...]
"""
import os
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

# name -> (cfg, nIter, dump iterations, arrays to drop (size), mesh file or None)
# name -> (cfg, nIter, dump iterations, iteration whose matrices are kept, mesh file or None)
CASES = {
    "naca0012_roe": ("naca0012_roe.cfg", 10, "0", 0, "mesh_NACA0012_inv.su2"),
    "naca0012_jst": ("naca0012_jst.cfg", 10, "0", None, "mesh_NACA0012_inv.su2"),
    "box_euler": ("box_euler.cfg", 6, "0,1,2", 2, None),
    "box_rans_sa": ("box_rans_sa.cfg", 6, "0,1,2", 2, None),
    "rect_ns": ("rect_ns.cfg", 6, "0,1,2", 2, None),
}
MATRIX_KEYS = ("_jacobian", "_ilu", "_invM")
KEEP_MATRICES = ("bc_jacobian", "_ilu", "_invM", "turb_bc_jacobian", "turb_src_jacobian", "turb_edge_jacobian", "turb_conv_jacobian")


def run_case(name, keep_all=False, out_dir=None):
    cfg, nit, dumps, keep_it, mesh = CASES[name]
    tmp = tempfile.mkdtemp(prefix="golden_")
    try:
        shutil.copy(os.path.join(ROOT, "tests", "cases", cfg), tmp)
        if mesh:
            shutil.copy(os.path.join(REF_MESH_DIR, mesh), tmp)
        exe = os.path.join(HERE, "_ref", "oracle_dump")
        subprocess.run([exe, cfg, "dump.bin", str(nit), dumps], cwd=tmp, check=True)
        d = load_dump(os.path.join(tmp, "dump.bin"))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    if not keep_all:
        for k in list(d.keys()):
            if any(k.endswith(m) for m in MATRIX_KEYS):
                keep = keep_it is not None and k.startswith("it%d_" % keep_it) and any(k.endswith(m) for m in KEEP_MATRICES)
                if name.startswith("naca") and not k.endswith("bc_jacobian"):
                    keep = False  # 36k blocks: one copy of the Jacobian is enough for the C1 mesh
                if not keep:
                    del d[k]
    out_dir = out_dir or os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, name + ".npz")
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in d.items()})
    print(name, "->", path, "%.2f MB" % (os.path.getsize(path) / 1e6))
    return path


if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES)
    for n in names:
        run_case(n)
