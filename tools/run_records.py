"""The long runs BASELINE.md §2 lists, on one GPU through the C ABI; writes profiles/r02_records_<case>.json.

  c1  Williamson 6 Rossby-Haurwitz 180x90, dt 300 s, 15 days   (BASELINE configs[0]; the CPU oracle runs it beside)
  c2  Williamson 5 mountain flow 720x360, dt 75 s, 15 days      (configs[1])
  c3  Jablonowski-Williamson baroclinic wave 360x180x32, dt 150 s, 10 days (configs[2]): min(phs) curve, growth onset
  c4  Held-Suarez 360x180x32, dt 150 s, N days with a cosine-bell tracer in a SECOND batch (configs[3])

Per day: tm, te, tpe (relative drift against day 0), extrema.  Nothing here imports oracle/ except --oracle (c1 / the
first steps of c4), which is the checker run beside the product, never inside it.
Usage: run_records.py CASE [DAYS] [--oracle]
"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from gmcore_b200.host import Host
from gmcore_b200.model import GmcoreGPU

CASES = {
    "c1": dict(nlon=180, nlat=90, nlev=1, dt=300.0, case="rossby_haurwitz_wave", days=15),
    "c2": dict(nlon=720, nlat=360, nlev=1, dt=75.0, case="mountain_zonal_flow", days=15),
    "c3": dict(nlon=360, nlat=180, nlev=32, dt=150.0, case="baroclinic_wave", days=10),
    "c4": dict(nlon=360, nlat=180, nlev=32, dt=150.0, case="held_suarez", days=100),
}


def cosine_bell(host, lon0=1.5 * np.pi, lat0=0.0, r0=1.0 / 3.0, h0=1.0):
    """The solid-rotation test's bell (tests/adv/solid_rotation_test_mod.F90:16-20) on every level, plus a 1e-3 floor."""
    q = np.zeros((1,) + host.shape("cell"))
    lon = host.vec["full_lon"][host.hx:host.hx + host.nlon]
    lat = host.vec["full_lat"][host.hy:host.hy + host.nlat]
    d = np.arccos(np.clip(np.sin(lat0) * np.sin(lat)[:, None] + np.cos(lat0) * np.cos(lat)[:, None] * np.cos(lon - lon0)[None, :], -1, 1))
    bell = np.where(d < r0, 0.5 * h0 * (1 + np.cos(np.pi * d / r0)), 0.0) + 1.0e-3
    host.interior(q[0], "cell")[...] = bell[None]
    return q


def main():
    name = sys.argv[1]
    c = dict(CASES[name])
    args = [a for a in sys.argv[2:] if not a.startswith("--")]
    if args:
        c["days"] = int(args[0])
    with_oracle = "--oracle" in sys.argv
    host = Host(c["nlon"], c["nlat"], c["nlev"], c["dt"], c["case"])
    ic = host.initial_condition()
    gpu = GmcoreGPU(host); gpu.set_state(ic)
    orc = None
    if with_oracle:
        from oracle.orc import Oracle
        orc = Oracle(host, fast=True); orc.set_state(ic)
    q0 = bell = None
    if host.cfg.ntracers:
        q0 = np.zeros((host.cfg.ntracers,) + host.shape("cell"))     # batch 0: the model's own qv-like tracer (zero here)
        gpu.set_tracers(q0)
        if orc: orc.set_tracers(q0)
    if name == "c4":
        bell = cosine_bell(host)
        for m in (gpu, orc):
            if m is not None:
                b = m.add_tracer_batch(1, c["dt"]); assert b == 1
                m.set_tracers(bell, b)
    gpu.prepare_run()
    if orc: orc.prepare_run()
    spd = int(round(86400.0 / c["dt"]))
    d0 = gpu.diagnose()
    area = np.asarray(host.vec["area_cell"][host.hy:host.hy + host.nlat])
    rec = {"case": name, "config": c, "hx": host.hx, "diag0": d0.tolist(), "days": [], "generator": "tools/run_records.py"}

    def tracer_mass(m):       # sum q * m * area of batch 1 (flux form: conserved to round-off)
        q = host.interior(m.get_tracers(1)[0], "cell"); mm = host.interior(m.field("m"), "cell")
        return float((q * mm * area[None, :, None]).sum()), float(q.min()), float(q.max())

    if name == "c4":
        rec["tracer0"] = tracer_mass(gpu)
    if orc is not None:
        # the checker beside the product: first steps, per-step relative L2 on the prognostic fields (+ the bell)
        nchk = 10 if name != "c1" else spd
        worst = 0.0
        names = [("u_lon", "lon"), ("v_lat", "lat")] + ([("gz", "cell")] if c["nlev"] == 1 else [("pt", "cell"), ("phs", "2d")])
        for s in range(nchk):
            gpu.run_steps(1); orc.run_steps(1)
            if s % max(1, nchk // 10) and s != nchk - 1:
                continue
            wind = np.hypot(np.linalg.norm(host.interior(orc.field("u_lon"), "lon")), np.linalg.norm(host.interior(orc.field("v_lat"), "lat")))
            for fn, loc in names:
                a, b = gpu.field(fn), orc.field(fn)
                if loc == "2d":
                    a = a.reshape(a.shape[-2:]); b = b.reshape(b.shape[-2:])
                a = host.interior(a, loc); b = host.interior(b, loc)
                den = wind if fn in ("u_lon", "v_lat") else np.linalg.norm(b)
                worst = max(worst, float(np.linalg.norm(a - b) / den))
            if name == "c4":
                a = host.interior(gpu.get_tracers(1)[0], "cell"); b = host.interior(orc.get_tracers(1)[0], "cell")
                worst = max(worst, float(np.linalg.norm(a - b) / np.linalg.norm(b)))
        rec["oracle_check"] = {"steps": nchk, "worst_rel_l2": worst, "bar": 1e-12 if nchk <= 10 else None}
        print("oracle check", rec["oracle_check"], flush=True)
        done = nchk
    else:
        done = 0
    t0 = time.perf_counter()
    for day in range(1, c["days"] + 1):
        n = day * spd - done
        if n <= 0:
            continue
        d = gpu.run_steps(n); done += n
        e = {"day": day, "tm_rel": float(d[-1, 0] / d0[0] - 1), "te_rel": float(d[-1, 1] / d0[1] - 1),
             "tpe_rel": float(d[-1, 3] / d0[3] - 1) if d0[3] else None,
             "tm_step_max": float(np.abs(np.diff(d[:, 0])).max() / d0[0]) if n > 1 else 0.0}
        if c["nlev"] > 1:
            phs = host.interior(gpu.field("phs").reshape(host.shape("2d")), "2d")
            e["phs_min"] = float(phs.min()); e["phs_max"] = float(phs.max())
            u = host.interior(gpu.field("u_lon"), "lon")
            e["u_max"] = float(np.abs(u).max())
            if name == "c4" and (day % 10 == 0 or day == c["days"]):
                zm = u.mean(axis=2)                       # zonal mean (lev, lat)
                e["zonal_mean_u_max"] = float(zm.max()); e["zonal_mean_u_min"] = float(zm.min())
                k, j = np.unravel_index(np.argmax(zm), zm.shape)
                e["jet"] = {"lev": int(k) + 1, "lat_deg": float(host.vec["full_lat_deg"][host.hy + j])}
                e["tracer"] = tracer_mass(gpu)
        else:
            gz = host.interior(gpu.field("gz"), "cell")
            e["gz_min"] = float(gz.min()); e["gz_max"] = float(gz.max())
        rec["days"].append(e)
        if day % 5 == 0 or day == c["days"]:
            print(json.dumps(e), flush=True)
    rec["wall_s"] = time.perf_counter() - t0
    rec["sdpd"] = c["days"] / rec["wall_s"] * 86400.0 if rec["wall_s"] > 0 else None
    if name == "c3":     # the wave must not have grown before day 4 (JW06: ps stays within ~1 hPa of 1000 hPa until then)
        d4 = [e for e in rec["days"] if e["day"] == 4]
        rec["no_growth_before_day4"] = bool(d4 and d4[0]["phs_min"] > 99800.0)
    if name == "c4":
        m0 = rec["tracer0"][0]; m1 = rec["days"][-1]["tracer"][0] if "tracer" in rec["days"][-1] else tracer_mass(gpu)[0]
        rec["tracer_mass_rel"] = float(m1 / m0 - 1)
        rec["tm_rel_per_step"] = rec["days"][-1]["tm_rel"] / (c["days"] * spd)
    os.makedirs(os.path.join(ROOT, "gpurun_out", "records"), exist_ok=True)
    out = os.path.join(ROOT, "gpurun_out", "records", f"r02_records_{name}.json")
    json.dump(rec, open(out, "w"), indent=1)
    print("wrote", out, "wall", round(rec["wall_s"], 1), "s", "SDPD", round(rec["sdpd"] or 0), flush=True)
    gpu.close()


if __name__ == "__main__":
    main()
