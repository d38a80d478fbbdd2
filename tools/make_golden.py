"""Generate tests/golden/*.json: regression fixtures from the CPU oracle (NOT from the Fortran
reference, which cannot be built here — see DESIGN.md "parity unpinned").  Each fixture holds the
per-step conserved integrals and L2 / max norms of the prognostic fields after the last step."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from test_oracle_cpu import run

CASES = {
    "swm_rh_64x33_wrfrk3": dict(nlon=64, nlat=33, nlev=1, dt=600.0, case="rossby_haurwitz_wave", nsteps=20, time_scheme="wrfrk3"),
    "swm_mz_64x33_pc2": dict(nlon=64, nlat=33, nlev=1, dt=600.0, case="mountain_zonal_flow", nsteps=20, time_scheme="pc2"),
    "bw_64x33x8_wrfrk3": dict(nlon=64, nlat=33, nlev=8, dt=600.0, case="baroclinic_wave", nsteps=8, time_scheme="wrfrk3"),
    "hs_48x24x8_wrfrk3": dict(nlon=48, nlat=24, nlev=8, dt=600.0, case="held_suarez", nsteps=8, time_scheme="wrfrk3"),
}
os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
for name, c in CASES.items():
    host, orc, d0, d = run(c["nlon"], c["nlat"], c["nlev"], c["dt"], c["case"], c["nsteps"], c["time_scheme"])
    fields = {}
    names = [("u_lon", "lon"), ("v_lat", "lat")] + ([("gz", "cell")] if c["nlev"] == 1 else [("pt", "cell"), ("phs", "2d")])
    for fname, loc in names:
        a = orc.field(fname)
        if loc == "2d":
            a = a.reshape(a.shape[-2:])
        a = host.interior(a, loc)
        fields[fname] = [loc, float(np.linalg.norm(a.ravel())), float(np.abs(a).max())]
    if c["nlev"] > 1:
        q = host.interior(orc.get_tracers()[0], "cell")
        fields_q = [float(np.linalg.norm(q.ravel())), float(np.abs(q).max())]
    else:
        fields_q = None
    json.dump({"generator": "tools/make_golden.py (CPU oracle, strict-FP build)", "config": c, "diag0": d0.tolist(),
               "diag": d.tolist(), "fields": fields, "tracer": fields_q}, open(os.path.join(ROOT, "tests", "golden", name + ".json"), "w"))
    print(name, "ok")
