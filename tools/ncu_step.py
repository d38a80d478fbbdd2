"""Short driver for ncu captures: prepare_run + N steps of one workload through the C ABI."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bench import WORKLOADS
from gmcore_b200.host import Host
from gmcore_b200.model import GmcoreGPU

name = sys.argv[1] if len(sys.argv) > 1 else "bw_1440x720x32"
nsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
wl = WORKLOADS[name]
host = Host(wl["nlon"], wl["nlat"], wl["nlev"], wl["dt"], wl["case"])
gpu = GmcoreGPU(host)
gpu.set_state(host.initial_condition())
if host.cfg.ntracers:
    from bench import bench_tracer
    gpu.set_tracers(bench_tracer(host))
gpu.prepare_run()
d = gpu.run_steps(nsteps)
print(name, "steps", nsteps, "launches", gpu.launches, "tm", d[-1, 0])
