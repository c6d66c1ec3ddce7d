"""Micro-benchmark of the edge kernel and the SpMV on the RANS box. usage: python tools/bench_edge.py N [reps]"""
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
ab = bench.algorithmic_bytes(3, z["nPoint"], z["nPointDomain"], z["nEdge"], z["nnz"], True)
ms = s.bench_edge_kernel(reps)
print("n=%d edge kernel %.3f ms  %.0f GB/s (algorithmic %.2f GB)" % (n, ms, ab["edge"] / ms / 1e6, ab["edge"] / 1e9))
ms = s.bench_spmv(0, reps)
print("n=%d spmv        %.3f ms  %.0f GB/s (algorithmic %.2f GB)" % (n, ms, ab["spmv"] / ms / 1e6, ab["spmv"] / 1e9))
s.close()
