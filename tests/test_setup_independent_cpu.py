"""The host set-up (libgmcore_host.so) against an independent restatement and reference-held tables.

Every GPU-vs-oracle parity test takes its mesh vectors, filter weights, vertical coordinate and initial state from
gmcore_b200/lib/libgmcore_host.so, on BOTH sides.  These tests close that gap: tests/setup_ref.py derives the same
quantities a second time from the Fortran (numpy r8 + mpmath 113-bit for the real(16) areas, nothing imported from
gmcore_b200/ or oracle/), and tests/golden/reference_tables.json holds the literals parsed out of the reference source
(tools/make_setup_golden.py).  Tolerance 1e-14 relative (areas: the two sides evaluate acos / atan2 in different
128-bit libraries; the r8 results agree to the last bit or two).
"""
import json
import os

import numpy as np
import pytest

import setup_ref as R
from gmcore_b200.host import Host

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-14

# BASELINE grids: C1 (180x90, dt 300), C3 (360x180, dt 150), C2 (720x360, dt 75), C5 (1440x720, dt 30)
GRIDS = [(180, 90, 300.0), (360, 180, 150.0), (720, 360, 75.0), (1440, 720, 30.0)]


def rel(a, b):
    """Largest element-wise relative difference; where the reference value is exactly 0 the other must be too."""
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    z = b == 0
    if (a[z] != 0).any():
        return float("inf")
    return float(np.max(np.abs(a[~z] - b[~z]) / np.abs(b[~z]))) if (~z).any() else 0.0


_CACHE = {}


def both(nlon, nlat, dt):
    key = (nlon, nlat, dt)
    if key not in _CACHE:
        _CACHE[key] = (Host(nlon, nlat, 1, dt, "rossby_haurwitz_wave"), R.mesh(nlon, nlat))
    return _CACHE[key]


@pytest.mark.parametrize("nlon,nlat,dt", GRIDS)
def test_mesh_vectors(nlon, nlat, dt):
    host, m = both(nlon, nlat, dt)
    hy, nh = host.hy, nlat - 1
    full = lambda n: host.vec[n][hy:hy + nlat]
    half = lambda n: host.vec[n][hy:hy + nh]
    worst = {}
    for n in ("full_lat", "full_cos_lat", "full_sin_lat", "area_cell", "area_lon", "area_lon_west", "area_lon_east",
              "area_lon_north", "area_lon_south", "de_lon", "le_lon", "full_f"):
        worst[n] = rel(full(n), m[n])
    for n in ("half_lat", "half_cos_lat", "half_sin_lat", "area_lat", "area_lat_west", "area_lat_east", "area_lat_north",
              "area_lat_south", "area_vtx", "de_lat", "le_lat", "half_f"):
        worst[n] = rel(half(n), m[n])
    sub = host.vec["area_subcell"].reshape(-1, 2)[hy:hy + nlat]
    worst["area_subcell"] = rel(sub, m["area_subcell"])
    # pole rows of full_tangent_wgt are never set by the reference: they keep the allocation value inf = huge(1.0_r8)
    # (mesh_mod.F90:656, loop bounds :420-423); the host does the same
    ftw = host.vec["full_tangent_wgt"].reshape(-1, 2)[hy:hy + nlat]
    assert (ftw[[0, -1]] == np.finfo(float).max).all()
    worst["full_tangent_wgt"] = rel(ftw[1:-1], m["full_tangent_wgt"][1:-1])
    worst["half_tangent_wgt"] = rel(host.vec["half_tangent_wgt"].reshape(-1, 2)[hy:hy + nh], m["half_tangent_wgt"])
    worst["half_lon"] = rel(host.vec["half_lon"][host.hx:host.hx + nlon], m["half_lon"])
    bad = {k: v for k, v in worst.items() if not v <= TOL}
    assert not bad, bad
    # exact identities the reference relies on (mesh_mod.F90:295-296,343-344,371-375)
    assert np.array_equal(full("area_lon_west"), full("area_lon_east"))
    assert half("area_lat_south")[0] == full("area_cell")[0] and half("area_lat_north")[-1] == full("area_cell")[-1]


@pytest.mark.parametrize("nlon,nlat,dt", GRIDS)
def test_filters_and_halo_width(nlon, nlat, dt):
    host, m = both(nlon, nlat, dt)
    hy, nh = host.hy, nlat - 1
    for which, kind in enumerate(("big_filter", "small_filter1", "small_filter2")):
        ngl, wl, ldl, nga, wa, lda = host.filters[which]
        rgl, rwl, rga, rwa = R.filters(m, dt, kind)
        assert np.array_equal(ngl[hy:hy + nlat], rgl), kind
        assert np.array_equal(nga[hy:hy + nh], rga), kind
        wl2 = wl.reshape(-1, ldl)[hy:hy + nlat]; wa2 = wa.reshape(-1, lda)[hy:hy + nh]
        worst = 0.0
        for j in range(nlat):
            if rwl[j] is not None:
                worst = max(worst, rel(wl2[j, :rgl[j]], rwl[j]))
                assert not wl2[j, rgl[j]:].any()
        for j in range(nh):
            if rwa[j] is not None:
                worst = max(worst, rel(wa2[j, :rga[j]], rwa[j]))
        # weights far in a Gaussian tail are ~1e-40 of the peak; compare relative to the row maximum instead there
        assert worst <= 1e-12, (kind, worst)
        if kind == "big_filter":
            assert host.hx == R.lon_halo_width(rga)
    # derived sizes quoted in SURVEY 8d / BASELINE.md
    expect_taps = {180: 47, 360: 85, 720: 161, 1440: 247}[nlon]
    assert int(host.filters[0][3].max()) == expect_taps


def test_wrf_l32_table_is_the_reference_table():
    """hyai / hybi of the 32-level template against the literals parsed from the reference source."""
    gold = json.load(open(os.path.join(HERE, "golden", "reference_tables.json")))
    host = Host(64, 33, 32, 600.0, "baroclinic_wave")
    # default-real literals: the Fortran compiler rounds them to binary32 first
    assert np.array_equal(host.vec["hyai"], np.array(gold["wrf_l32"]["hyai_r4"]))
    assert np.array_equal(host.vec["hybi"], np.array(gold["wrf_l32"]["hybi_r4"]))
    assert np.array_equal(np.float32(np.array([float(x) for x in gold["wrf_l32"]["hyai_text"]])).astype(float), host.vec["hyai"])
    full_lev, half_lev = R.vert_levels(host.vec["hyai"], host.vec["hybi"])
    assert rel(host.vec["half_lev"], half_lev) <= TOL
    assert rel(host.vec["full_lev"][2:2 + 32], full_lev) <= TOL
    e = gold["earth"]
    c = host.cfg
    assert (c.radius, c.g, c.Rd, c.cpd, c.p0) == (e["radius"], e["g"], e["Rd"], e["cpd"], e["p0"])
    assert c.omega == 2 * R.PI / 86400.0


@pytest.mark.parametrize("case,nlon,nlat,dt", [("rossby_haurwitz_wave", 180, 90, 300.0), ("mountain_zonal_flow", 720, 360, 75.0)])
def test_swm_initial_conditions(case, nlon, nlat, dt):
    host = Host(nlon, nlat, 1, dt, case)
    _, m = both(nlon, nlat, dt)
    ic = host.initial_condition()
    ref = R.ic_rossby_haurwitz(m) if case == "rossby_haurwitz_wave" else R.ic_mountain_zonal_flow(m)
    scale = lambda a: np.abs(a).max()
    for name, key, loc in (("u", "u_lon", "lon"), ("v", "v_lat", "lat"), ("gz", "pt_or_gz", "cell")):
        a = host.interior(ic[key], loc)[0]
        assert np.abs(a - ref[name]).max() <= 4e-15 * max(scale(ref[name]), 1e-300) + 1e-300, name
    assert np.abs(host.interior(ic["gzs"], "2d") - ref["gzs"]).max() <= 4e-15 * max(scale(ref["gzs"]), 1.0)


@pytest.mark.parametrize("nlon,nlat,dt", [(360, 180, 150.0)])
def test_baroclinic_wave_initial_condition(nlon, nlat, dt):
    host = Host(nlon, nlat, 32, dt, "baroclinic_wave")
    _, m = both(nlon, nlat, dt)
    ic = host.initial_condition()
    ref = R.ic_baroclinic_wave(m, host.vec["hyai"], host.vec["hybi"])
    u = host.interior(ic["u_lon"], "lon"); pt = host.interior(ic["pt_or_gz"], "cell")
    assert np.abs(u - ref["u"]).max() <= 4e-15 * np.abs(ref["u"]).max()
    assert np.abs(pt - ref["pt"]).max() <= 4e-15 * np.abs(ref["pt"]).max()
    assert not host.interior(ic["v_lat"], "lat").any()
    assert np.array_equal(host.interior(ic["phs"], "2d"), ref["phs"])
    assert np.abs(host.interior(ic["gzs"], "2d") - ref["gzs"]).max() <= 4e-15 * np.abs(ref["gzs"]).max()
