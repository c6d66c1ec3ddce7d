"""Where the step time goes that no profile category covers (SU2B200_PROFILE_GAPS=1). usage: python tools/profile_gaps.py [n]"""
import os, sys
os.environ["SU2B200_PROFILE_GAPS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from su2_b200.solver import FlowSolver
from su2_b200.config import parse_cfg, flow_config_from_cfg, mesh_from_cfg
n = int(sys.argv[1]) if len(sys.argv) > 1 else 185
opts = parse_cfg(bench.make_cfg("RANS", [n, n, n], 1))
mesh = mesh_from_cfg(opts); s = FlowSolver(0); mesh.upload(s)
s.configure(**flow_config_from_cfg(opts, 3)); s.allocate(); mesh.close()
s.iterate(8)
s.timer_start(); s.iterate(4); t = s.timer_stop() / 4
s.profile_enable(True)
s.timer_start(); s.iterate(4); tp = s.timer_stop() / 4
prof = s.profile_report()
s.profile_enable(False)
print("step %.2f ms unprofiled, %.2f ms profiled; categories sum %.2f ms" % (t, tp, sum(v[0] for v in prof.values()) / 4))
s.close()
