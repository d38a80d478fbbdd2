"""Per-kernel CUDA-event timing of K steps of one workload (gmc_profile): name count ms/step mean_ms."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bench import WORKLOADS
from gmcore_b200.host import Host
from gmcore_b200.model import GmcoreGPU

name = sys.argv[1] if len(sys.argv) > 1 else "bw_1440x720x32"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 5
wl = WORKLOADS[name]
host = Host(wl["nlon"], wl["nlat"], wl["nlev"], wl["dt"], wl["case"])
gpu = GmcoreGPU(host)
gpu.set_state(host.initial_condition())
if host.cfg.ntracers:
    from bench import bench_tracer
    gpu.set_tracers(bench_tracer(host))
gpu.prepare_run()
gpu.run_steps(3)
gpu.synchronize()
gpu.timer_start(); d = gpu.run_steps(K); ms = gpu.timer_stop_ms()
gpu.profile(True); gpu.run_steps(K); rep = gpu.profile_report(); gpu.profile(False)
print(f"{name} lib={os.environ.get('GMCORE_B200_LIB', 'default')} ms/step={ms / K:.3f} SDPD={wl['dt'] * K / (ms / 1e3):.1f} tm={d[-1, 0]:.15e} te={d[-1, 1]:.15e}")
for k, v in sorted(rep.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k:18s} n/step={v[0] / K:6.1f} ms/step={v[1] / K:8.4f} mean_ms={v[1] / v[0]:8.4f}")
