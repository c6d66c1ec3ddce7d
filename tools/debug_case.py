"""Print the per-stage relative errors of one golden case (run on the GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.parity_util import make_solver, run_first_iteration, load_golden, ROOT
case = sys.argv[1]
gold = load_golden(os.path.join(ROOT, "tests", "golden", case + ".npz"))
its = sorted({int(k[2:k.index("_")]) for k in gold if k.startswith("it") and k.endswith("_pre_solution")})
for it in its:
    s = make_solver(gold)
    rep = {"_keep": {}}
    run_first_iteration(s, gold, it, rep)
    keep = rep.pop("_keep")
    for name in sys.argv[2:]:
        if name in keep:
            got, ref = keep[name]
            got = got.reshape(ref.shape) if got.size == ref.size else got
            d = np.abs(got - ref)
            idx = np.argsort(d.ravel())[::-1][:8]
            print("  top diffs of", name, "shape", ref.shape)
            for k in idx:
                u = np.unravel_index(k, ref.shape)
                print("    ", u, "got %.15e ref %.15e" % (got[u], ref[u]))
    print(case, "it", it)
    for k, v in rep.items():
        print("   %-18s %s" % (k, ("%.3e" % v) if isinstance(v, float) else v))
    s.close()
