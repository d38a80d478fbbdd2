"""One instance of the CPU restatement (oracle, -O3 OpenMP build) for the timed CPU baseline.
Usage: cpu_worker.py WORKLOAD BUDGET_S [MIN_STEPS [MAX_STEPS]]  ->  prints "steps seconds rss_mb threads".
Threads come from OMP_NUM_THREADS (the level loops of the restatement are OpenMP-parallel in this build)."""
import os, sys, time, resource
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bench import WORKLOADS
from gmcore_b200.host import Host
from oracle.orc import Oracle

name, budget = sys.argv[1], float(sys.argv[2])
min_steps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
max_steps = int(sys.argv[4]) if len(sys.argv) > 4 else 200
wl = WORKLOADS[name]
host = Host(wl["nlon"], wl["nlat"], wl["nlev"], wl["dt"], wl["case"])
orc = Oracle(host, fast=os.environ.get("ORC_FAST", "1") == "1")   # ORC_FAST=0: the scalar -O2 parity build
orc.set_state(host.initial_condition())
if host.cfg.ntracers:
    from bench import bench_tracer
    orc.set_tracers(bench_tracer(host))
orc.prepare_run()
t0 = time.perf_counter(); orc.run_steps(1); t1 = time.perf_counter() - t0      # warm-up step (first touch of the work arrays)
n = int(max(min_steps, min(max_steps, budget / max(t1, 1e-6))))
t0 = time.perf_counter(); orc.run_steps(n); el = time.perf_counter() - t0
print(n, el, resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1024.0, orc.lib.orc_num_threads())
