"""Micro-benchmark of the ILU0 build / apply kernels on the RANS box (run on the GPU box).
usage: python tools/bench_prec.py N [reps]   (SU2B200_ILU_APPLY=0|1|2, SU2B200_ILU_BUILD=0|1 select the schedules)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from su2_b200.solver import FlowSolver
from su2_b200.config import parse_cfg, flow_config_from_cfg, mesh_from_cfg
n = int(sys.argv[1]); reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
opts = parse_cfg(bench.make_cfg("RANS", [n, n, n], 1))
tets = bool(int(os.environ.get("TETS", "0")))
mesh = mesh_from_cfg(opts, tets=tets); s = FlowSolver(0); mesh.upload(s)
s.configure(**flow_config_from_cfg(opts, 3)); s.allocate()
s.iterate(2)
z = mesh.sizes()
for which, nv in ((0, 5), (1, 1)):
    b, a = s.bench_prec(which, reps=reps)
    nnz, Pd = z["nnz"], z["nPointDomain"]
    bytes_apply = nnz * (8 * nv * nv + 4) + Pd * 8 * nv * nv + Pd * 3 * 8 * nv
    bytes_build = 2 * nnz * 8 * nv * nv + Pd * 8 * nv * nv
    print("n=%d sys=%d nVar=%d build %.3f ms (%.0f GB/s)  apply %.3f ms (%.0f GB/s)  mode=%s" % (
        n, which, nv, b, bytes_build / b / 1e6, a, bytes_apply / a / 1e6,
        "apply=%s build=%s" % (os.environ.get("SU2B200_ILU_APPLY", "2"), os.environ.get("SU2B200_ILU_BUILD", "0"))))
s.close()
