"""One single-thread replica of the CPU restatement (oracle, -O3 build) for the timed CPU baseline.
Usage: cpu_worker.py WORKLOAD BUDGET_S  ->  prints "steps seconds rss_mb"."""
import os, sys, time, resource
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bench import WORKLOADS
from gmcore_b200.host import Host
from oracle.orc import Oracle

name, budget = sys.argv[1], float(sys.argv[2])
wl = WORKLOADS[name]
host = Host(wl["nlon"], wl["nlat"], wl["nlev"], wl["dt"], wl["case"])
orc = Oracle(host, fast=os.environ.get("ORC_FAST", "1") == "1")   # ORC_FAST=0: the portable -O2 build
orc.set_state(host.initial_condition())
if host.cfg.ntracers:
    orc.set_tracers(np.zeros((host.cfg.ntracers,) + host.shape("cell")))
orc.prepare_run()
t0 = time.perf_counter(); orc.run_steps(1); t1 = time.perf_counter() - t0
n = int(max(1, min(200, budget / max(t1, 1e-6))))
t0 = time.perf_counter(); orc.run_steps(n); el = time.perf_counter() - t0
print(n, el, resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1024.0)
