"""CPU checks of the stand-alone host set-up against the reference's own invariants:
debug_check_areas (src/utils/debug_mod.F90:33-111) and the filter construction
(src/dynamics/filter_types_mod.F90:49-147)."""
import numpy as np
import pytest

from gmcore_b200.host import Host


@pytest.fixture(scope="module", params=[(180, 90, 1, 300.0), (360, 181, 1, 150.0), (64, 33, 8, 600.0)])
def host(request):
    nlon, nlat, nlev, dt = request.param
    return Host(nlon, nlat, nlev, dt, "rossby_haurwitz_wave" if nlev == 1 else "baroclinic_wave")


def F(host, name):   # full-lat vector, interior rows 1..nlat
    return host.vec[name][host.hy:host.hy + host.nlat]


def H(host, name):   # half-lat vector, rows 1..nlat-1
    return host.vec[name][host.hy:host.hy + host.nlat - 1]


def test_debug_check_areas(host):
    total = 4 * np.pi * host.cfg.radius ** 2
    nlon = host.nlon
    assert abs(F(host, "area_cell").sum() * nlon - total) / total < 1e-12
    assert abs(H(host, "area_vtx").sum() * nlon - total) / total < 1e-12
    sub = host.vec["area_subcell"].reshape(-1, 2)[host.hy:host.hy + host.nlat]       # (j, 2)
    assert abs(sub.sum() * nlon * 2 - total) / total < 1e-12
    assert np.all(np.abs(F(host, "area_cell") - 2 * sub.sum(axis=1)) / F(host, "area_cell") < 1e-12)
    vtx = H(host, "area_vtx")
    assert np.all(np.abs(vtx - 2 * (sub[:-1, 1] + sub[1:, 0])) / vtx < 1e-12)
    al = F(host, "area_lon")[1:-1]
    assert np.all(np.abs(F(host, "area_lon_north")[1:-1] + F(host, "area_lon_south")[1:-1] - al) / al < 1e-12)
    at = H(host, "area_lat")
    assert np.all(np.abs(H(host, "area_lat_west") + H(host, "area_lat_east") - at) / at < 1e-11)
    edge = (al.sum() + at.sum()) * nlon
    assert abs(edge - total) / total < 1e-9


def test_geometry_identities(host):
    """le_lon = 2 area_lon / de_lon, de_lat = 2 area_lat / le_lat (mesh_mod.F90:381,394); west == east."""
    al, de, le = F(host, "area_lon")[1:-1], F(host, "de_lon")[1:-1], F(host, "le_lon")[1:-1]
    np.testing.assert_allclose(le, 2 * al / de, rtol=1e-14)
    at, de2, le2 = H(host, "area_lat"), H(host, "de_lat"), H(host, "le_lat")
    np.testing.assert_allclose(de2, 2 * at / le2, rtol=1e-14)
    assert np.array_equal(F(host, "area_lon_west"), F(host, "area_lon_east"))
    assert np.array_equal(H(host, "area_lat_west"), H(host, "area_lat_east"))
    # equatorial symmetry of the latitudes (mesh_mod.F90:226-237)
    assert np.array_equal(F(host, "full_lat"), -F(host, "full_lat")[::-1])
    assert np.array_equal(H(host, "half_lat"), -H(host, "half_lat")[::-1])


def test_filters(host):
    """Odd widths, normalised Gaussian weights, n = 0 on pole rows, big filter n >= 3 elsewhere,
    lon halo width = max((max n - 1)/2, 2) (process_mod.F90:196-268)."""
    big_n = None
    for which, (ngl, wl, ldl, nga, wa, lda) in enumerate(host.filters):
        n = ngl[host.hy:host.hy + host.nlat]
        w = wl.reshape(-1, ldl)[host.hy:host.hy + host.nlat]
        assert n[0] <= 1 and n[-1] <= 1
        for j in range(1, host.nlat - 1):
            if n[j] > 1:
                assert n[j] % 2 == 1
                assert abs(w[j, :n[j]].sum() - 1) < 1e-14
                np.testing.assert_allclose(w[j, :n[j]], w[j, :n[j]][::-1], rtol=1e-13)
        if which == 0:
            assert np.all(n[1:-1] >= 3)
            big_n = max(n.max(), nga.max())
    assert host.hx == max((big_n - 1) // 2, 2)


def test_vertical_coordinate():
    h = Host(64, 33, 32, 600.0, "baroclinic_wave")
    hyai, hybi = h.vec["hyai"], h.vec["hybi"]
    assert hyai.shape == (33,) and hybi[0] == 0 and hybi[-1] == 1 and hyai[-1] == 0
    assert np.all(np.diff(hyai + hybi) > 0)
    # wrf_l32 table entries are single-precision literals widened to r8
    assert np.array_equal(hyai, hyai.astype(np.float32).astype(np.float64))
    np.testing.assert_allclose(h.vec["half_lev"], hyai + hybi, rtol=0, atol=0)


def test_coarse_pole_and_thuburn09_mesh():
    """coarse_pole_mul /= 0 (mesh_mod.F90:195-202): dlat grows towards the poles, the rows still tile the sphere
    and stay mirror-symmetric.  tangent_wgt_scheme = 'thuburn09' (mesh_mod.F90:428-437): sub-cell area ratios
    replace the constant 1/4; on every row the two weights still sum to the classic pair's sum scaled by the
    sub-cell split, and the oracle conserves mass on the stretched mesh."""
    from oracle.orc import Oracle
    nlon, nlat = 64, 33
    h0 = Host(nlon, nlat, 8, 600.0, "baroclinic_wave")
    h = Host(nlon, nlat, 8, 600.0, "baroclinic_wave", coarse_pole_mul=2.0, coarse_pole_decay=100.0,
             tangent_wgt_scheme="thuburn09")
    hy = h.hy
    fl = h.vec["full_lat"][hy:hy + nlat]; hl = h.vec["half_lat"][hy:hy + nlat - 1]
    d = np.diff(fl)
    assert d[0] / d[nlat // 2] > 1.5 and np.all(d > 0)
    assert np.allclose(fl, -fl[::-1], rtol=0, atol=1e-15) and np.allclose(hl, -hl[::-1], rtol=0, atol=1e-15)
    a = h.vec["area_cell"][hy:hy + nlat]
    assert abs(a.sum() * nlon / (4 * np.pi * h.cfg.radius ** 2) - 1) < 1e-14
    ftw, ftw0 = h.vec["full_tangent_wgt"].reshape(-1, 2), h0.vec["full_tangent_wgt"].reshape(-1, 2)
    j = slice(hy + 1, hy + nlat - 1)
    assert np.all(np.isfinite(ftw[j])) and np.all(ftw[j] > 0) and not np.allclose(ftw[j], ftw0[j])
    with pytest.raises(ValueError):
        Host(nlon, nlat, 8, 600.0, "baroclinic_wave", tangent_wgt_scheme="bogus")
    o = Oracle(h); o.set_state(h.initial_condition()); o.prepare_run()
    dg = o.run_steps(5)
    assert abs(dg[-1, 0] / dg[0, 0] - 1) < 1e-13
    assert abs(dg[-1, 1] / dg[0, 1] - 1) < 1e-5
