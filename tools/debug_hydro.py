"""Developer aid: print the per-field GPU-vs-oracle relative L2 for a small hydrostatic run."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from util import make_pair, rel_l2, smooth_tracer, PROGNOSTIC_HYDRO
from test_hydro_gpu import DERIVED, WORK, field_errors

case = sys.argv[1] if len(sys.argv) > 1 else "baroclinic_wave"
nlon, nlat, nlev, dt = (int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])) if len(sys.argv) > 5 else (64, 33, 8, 600.0)
scheme = sys.argv[6] if len(sys.argv) > 6 else "wrfrk3"
host, ic, orc, gpu = make_pair(nlon, nlat, nlev, dt, case, time_scheme=scheme)
q = smooth_tracer(host, host.cfg.ntracers)
orc.set_tracers(q); gpu.set_tracers(q)
orc.prepare_run(); gpu.prepare_run()
def show(tag, fields):
    e = field_errors(host, orc, gpu, fields)
    print(tag, " ".join(f"{k}={v:.1e}" for k, v in e.items()), flush=True)
show("prepare", DERIVED)
print("diag gpu", gpu.diagnose()); print("diag orc", orc.diagnose())
for s in range(3):
    dg = gpu.run_steps(1); do = orc.run_steps(1)
    show(f"step{s} prog", PROGNOSTIC_HYDRO)
    show(f"step{s} derived", DERIVED)
    show(f"step{s} work", [w for w in WORK if w[0] not in ("du", "dv", "dpt")])
    print(f"step{s} tracer", rel_l2(host.interior(gpu.get_tracers()[0], "cell"), host.interior(orc.get_tracers()[0], "cell")))
    print("diag rel", np.abs(dg / np.where(do == 0, 1, do) - 1))
