"""Bitwise check of the pure-pressure-level table (DMesh.lev_tab): the same run with the table on and off
(GMCORE_NO_LEVTAB=1 is read when the mesh is set) must give identical states."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmcore_b200.host import Host          # noqa: E402
from gmcore_b200.model import GmcoreGPU    # noqa: E402

host = Host(180, 91, 32, 300.0, "baroclinic_wave")
ic = host.initial_condition()
out = []
for off in (False, True):
    if off:
        os.environ["GMCORE_NO_LEVTAB"] = "1"
    else:
        os.environ.pop("GMCORE_NO_LEVTAB", None)
    gpu = GmcoreGPU(host)
    gpu.set_state(ic)
    gpu.set_tracers(np.ones((host.cfg.ntracers,) + host.shape("cell")))
    gpu.prepare_run()
    gpu.run_steps(5)
    out.append(gpu.download_state())
    gpu.close()
same = True
for key in out[0]:
    a, b = out[0][key], out[1][key]
    a = np.asarray(a); b = np.asarray(b)
    eq = np.array_equal(a, b)
    same &= eq
    print("equal" if eq else f"DIFF max {np.abs(a - b).max():.3e}")
print("levtab bitwise identical:", same)
sys.exit(0 if same else 1)
