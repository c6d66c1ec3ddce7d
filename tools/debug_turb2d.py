"""Debug helper (GPU box): where does the 2-D SA Jacobian differ from the reference's? usage: python tools/debug_turb2d.py [case] [it]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.parity_util import make_solver, run_first_iteration, load_case
case = sys.argv[1] if len(sys.argv) > 1 else "rect_rans_sa"
it = int(sys.argv[2]) if len(sys.argv) > 2 else 2
gold = load_case(case)
s = make_solver(gold)
rep = {}
run_first_iteration(s, gold, it, rep)
pre = "it%d_turb_" % it
s.Turb_Preprocessing()
s.set("TURB_SOLUTION_OLD", s.get("TURB_SOLUTION"))
s.Turb_SetTime_Step()
s.Turb_Space_Integration()
J = s.get("TURB_JACOBIAN")
R = s.get("TURB_LINSYSRES")
rp = np.asarray(gold["row_ptr"]).astype(np.int64); ci = np.asarray(gold["col_ind"]).astype(np.int64); dp = np.asarray(gold["dia_ptr"]).astype(np.int64)
P = len(rp) - 1
for key in ("conv_jacobian", "edge_jacobian", "src_jacobian", "bc_jacobian"):
    if pre + key in gold:
        G = np.asarray(gold[pre + key]).ravel()
        d = np.abs(J - G)
        print(key, "max abs diff %.3e (scale %.3e)" % (d.max(), np.abs(G).max()))
G = np.asarray(gold[pre + "bc_jacobian"]).ravel()
d = np.abs(J - G)
isdiag = np.zeros(len(J), bool); isdiag[dp] = True
print("diag max diff %.3e  offdiag max diff %.3e" % (d[isdiag].max(), d[~isdiag].max()))
row_of = np.repeat(np.arange(P), np.diff(rp))
worst = np.argsort(d)[-8:]
pb = np.asarray(gold["physical_boundary"]).ravel()
for k in worst:
    print("entry", k, "row", row_of[k], "col", ci[k], "mine %.12e ref %.12e" % (J[k], G[k]), "boundary row" if pb[row_of[k]] else "interior", "diag" if isdiag[k] else "off")
print("res diff", np.abs(R - np.asarray(gold[pre + "bc_linsysres"]).ravel()).max())
s.Turb_PrepareImplicitIteration()
iters, res = s.Turb_Solve()
print("iters", iters, "sol relerr", np.abs(s.get("TURB_LINSYSSOL") - np.asarray(gold[pre + "linsyssol"]).ravel()).max() / np.abs(gold[pre + "linsyssol"]).max())
s.close()
