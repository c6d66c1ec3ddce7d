"""Per-step device times of the bench workload (diagnostic): python tools/step_times.py [n] [steps]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from su2_b200.config import parse_cfg, flow_config_from_cfg, mesh_from_cfg
from su2_b200.solver import FlowSolver
import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 185
K = int(sys.argv[2]) if len(sys.argv) > 2 else 30
text = bench.make_cfg("RANS", [n, n, n], 1)
opts = parse_cfg(text)
mesh = mesh_from_cfg(opts, rank=0, nranks=1)
s = FlowSolver(0)
mesh.upload(s)
s.configure(**flow_config_from_cfg(opts, 3))
s.allocate()
mesh.close()
for _ in range(5):
    s.iterate(1)
ts = []
for _ in range(K):
    s.timer_start()
    r = s.iterate(1)
    ts.append(s.timer_stop())
print("single steps:", " ".join("%.1f" % t for t in ts))
s.timer_start()
s.iterate(K)
print("one call of %d: %.2f ms/step" % (K, s.timer_stop() / K))
s.timer_start()
s.iterate(K)
print("again: %.2f ms/step" % (s.timer_stop() / K))
s.close()
