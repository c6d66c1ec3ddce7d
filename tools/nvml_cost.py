import os, sys, time, threading
sys.path.insert(0, "/root/repo")
import bench
from su2_b200.solver import FlowSolver
from su2_b200.config import parse_cfg, flow_config_from_cfg, mesh_from_cfg
import pynvml as nv
n = 129
opts = parse_cfg(bench.make_cfg("RANS", [n, n, n], 1))
mesh = mesh_from_cfg(opts); s = FlowSolver(0); mesh.upload(s)
s.configure(**flow_config_from_cfg(opts, 3)); s.allocate(); s.iterate(3)
nv.nvmlInit(); h = nv.nvmlDeviceGetHandleByIndex(0)
def timed(K):
    s.timer_start(); s.iterate(K); return s.timer_stop() / K
print("quiet ms/step", timed(10), timed(10))
stop = threading.Event(); cost = {}
def sampler(which, period):
    while not stop.is_set():
        t0 = time.perf_counter()
        if which == "clock": nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        elif which == "reasons": nv.nvmlDeviceGetCurrentClocksEventReasons(h)
        elif which == "power": nv.nvmlDeviceGetPowerUsage(h)
        elif which == "smi":
            import subprocess
            subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks_event_reasons.active", "--format=csv,noheader"], stdout=subprocess.PIPE)
        cost.setdefault(which, []).append(time.perf_counter() - t0)
        stop.wait(period)
for which, period in (("clock", 0.05), ("reasons", 0.05), ("power", 0.05), ("clock", 0.005), ("smi", 0.2)):
    stop.clear(); cost.clear()
    th = threading.Thread(target=sampler, args=(which, period)); th.start()
    t = timed(20)
    stop.set(); th.join()
    c = cost[which]
    print("%-8s period %.3f: ms/step %.2f, %d calls, mean call %.2f ms" % (which, period, t, len(c), 1e3 * sum(c) / len(c)))
