"""Developer aid: compare an N-band group with the single-band run field by field, row by row."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from util import make_pair, smooth_tracer
from gmcore_b200.model import GmcoreBandGroup, FIELD_LOC, round_robin

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 2
scheme = sys.argv[2] if len(sys.argv) > 2 else "euler"
host, ic, orc, one = make_pair(64, 33, 8, 600.0, "baroclinic_wave", time_scheme=scheme)
q = smooth_tracer(host, 1); one.set_tracers(q)
grp = GmcoreBandGroup(host, nb); grp.set_state(ic); grp.set_tracers(q)
print("bands", round_robin(33, nb))
one.prepare_run(); grp.prepare_run()
names = ["u_f", "v_f", "phs_f", "m", "m_lev", "ph_exn_lev", "t", "gz_lev", "mfx_lon", "mfy_lat", "u_lat", "v_lon", "ke", "pv", "pv_lon", "pv_lat", "div",
         "cflx", "cfly", "divx", "divy", "we_lev", "cflz", "dphs", "qx", "qy", "ptf_lon", "ptf_lat", "dpt", "du", "dv", "pt", "phs", "u_lon", "v_lat"]
def cmp(tag):
    for n in names:
        loc = FIELD_LOC[n]
        a, b = grp.field(n), one.field(n)
        if loc == "2d": a = a.reshape(a.shape[-2:]); b = b.reshape(b.shape[-2:])
        a, b = host.interior(a, loc), host.interior(b, loc)
        d = np.abs(a - b)
        if d.ndim == 3: d = d.max(axis=(0, 2))
        else: d = d.max(axis=1)
        bad = np.nonzero(d > 0)[0]
        print(tag, n, "ok" if len(bad) == 0 else f"rows(0-based) {bad.tolist()[:12]} max {d.max():.2e}")
cmp("prep")
one.run_steps(1); grp.run_steps(1)
cmp("step1")
# band-local arrays incl. halo rows vs the single-band run
hy = host.hy
for b in grp.bands:
    jb, je = b.cfg.lat_ibeg, b.cfg.lat_iend
    for n in ["pt", "phs_f", "m", "gz_lev", "ph_exn_lev", "t", "v_lat", "u_f", "v_f", "phs", "we_lev", "ptf_lat", "ptf_lon", "cflz", "qx", "qy", "dphs"]:
        loc = FIELD_LOC[n]
        a, g = b.field(n), one.field(n)
        if loc == "2d": a = a.reshape((1,) + a.shape[-2:]); g = g.reshape((1,) + g.shape[-2:])
        elif not loc.endswith("+lev"): a = a[2:2 + host.nlev]; g = g[2:2 + host.nlev]
        half = loc.split("+")[0] in ("lat", "vtx")
        out = []
        for j in range(max(1, jb - 2), min(host.nlat - (1 if half else 0), je + 2) + 1):
            rb = j - jb + hy
            if rb < 0 or rb >= a.shape[1]: continue
            d = np.abs(a[:, rb, host.hx:host.hx + host.nlon] - g[:, j - 1 + hy, host.hx:host.hx + host.nlon]).max()
            if d > 0: out.append((j, float(d)))
        print("band", b.cfg.rank, n, out)
