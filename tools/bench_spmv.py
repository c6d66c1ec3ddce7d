"""Micro-benchmark of the block SpMV on the RANS box (run on the GPU box). usage: python tools/bench_spmv.py N [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from su2_b200.solver import FlowSolver
from su2_b200.config import parse_cfg, flow_config_from_cfg, mesh_from_cfg
n = int(sys.argv[1]); reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
opts = parse_cfg(bench.make_cfg("RANS", [n, n, n], 1))
mesh = mesh_from_cfg(opts); s = FlowSolver(0); mesh.upload(s)
s.configure(**flow_config_from_cfg(opts, 3)); s.allocate()
s.iterate(2)
z = mesh.sizes()
nnz, Pd, P, nv = z["nnz"], z["nPointDomain"], z["nPoint"], 5
t = s.bench_spmv(0, reps=reps)
byt = nnz * (8 * nv * nv + 4) + Pd * (4 + 8 * nv) + P * 8 * nv
print("n=%d spmv %.3f ms (%.0f GB/s)" % (n, t, byt / t / 1e6))
s.close()
