"""Independent re-derivation of the host set-up, for tests only.

The parity tests feed the CUDA path and the oracle from ONE host library (gmcore_b200/lib/libgmcore_host.so:
mesh geometry, filter weights, vertical coordinate, initial conditions).  A mis-transcribed coefficient there
would pass every GPU-vs-oracle test.  This module states the same set-up a second time, straight from the
reference Fortran, in a different language and arithmetic (numpy r8 + mpmath at 113 bits for the real(16)
spherical areas), with no import of gmcore_b200 or oracle.  tests/test_setup_independent_cpu.py compares the
two on the BASELINE grids.

Reference (GMCORE source tree):
  src/utils/const_mod.F90:12-66                        constants
  src/dynamics/mesh_mod.F90:134-447                    mesh_init_global
  src/utils/sphere_geometry_mod.F90:110-146,302-427    cartesian transforms, calc_area, small-arc lune, sphere angle
  src/dynamics/filter_types_mod.F90:31-147             gaussian_weight, filter_init
  src/utils/math_mod.F90:146-159                       exp_two_values
  src/utils/process_mod.F90:206-211                    lon_halo_width from big_filter%ngrid_lat
  src/dynamics/vert_coord/hybrid_coord_mod.F90:124-137 full_lev / half_lev from hyai / hybi
  src/tests/swm/rossby_haurwitz_wave_test_mod.F90:31-87, mountain_zonal_flow_test_mod.F90:24-81
  src/tests/baroclinic_wave_test_mod.F90:28-170
Single-precision literals of the reference are rounded through float32 here as well.
"""
import math

import mpmath as mp
import numpy as np

f32 = lambda x: float(np.float32(x))

PI = math.atan(1.0) * 4.0
PI2 = PI * 2
PI05 = PI * 0.5
DEG = 180.0 / PI
OMEGA = 2 * PI / 86400.0
RADIUS = 6.37122e6
G = 9.80616
RD = 287.04
CPD = 1004.0
P0 = 1.0e5
RD_O_CPD = RD / CPD

# namelist defaults (namelist_mod.F90:53,89-95): default-real literals widened to r8
MAX_WAVE_SPEED = 300.0
MAX_CFL = f32(0.5)
FILTER_A, FILTER_B, FILTER_C, FILTER_D, FILTER_E = f32(1.0), f32(0.4), f32(0.2), f32(0.375), f32(0.16)
PTOP = 2.194e2

mp.mp.prec = 113          # IEEE binary128, the reference's real(16)
_Q = mp.mpf
_QPI = _Q(PI)             # the r8 constant pi widened (sphere_geometry_mod uses const_mod's pi)
_QPI2 = _Q(PI2)
_QR = _Q(RADIUS)


def _cart(lon, lat):
    """cartesian_transform_1_r16: cos / sin evaluated in r8, products in r16 (sphere_geometry_mod.F90:110-122)."""
    cos_lat = _Q(math.cos(lat))
    return (_QR * cos_lat * _Q(math.cos(lon)), _QR * cos_lat * _Q(math.sin(lon)), _QR * _Q(math.sin(lat)))


def _cross(a, b):
    return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])


def _dot(a, b):
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def _norm(x):
    n = mp.sqrt(_dot(x, x))
    return tuple(c / n for c in x) if n != 0 else x


def _angle(a, b, c):
    """calc_sphere_angle_1 (sphere_geometry_mod.F90:410-427)."""
    nab = _norm(_cross(a, b)); nbc = _norm(_cross(b, c))
    res = mp.acos(-max(min(_dot(nab, nbc), _Q(1)), _Q(-1)))
    if _dot(_cross(nab, nbc), a) < 0:
        res = _QPI2 - res
    return res


def _area(p):
    """calc_area: spherical excess of the polygon p (sphere_geometry_mod.F90:302-326)."""
    n = len(p)
    s = _Q(0)
    for i in range(n):
        s = s + _angle(p[i - 1], p[i], p[(i + 1) % n])
    return _QR ** 2 * (s - (n - 2) * _QPI)


def _area_small_arc(p):
    """calc_area_with_last_small_arc (sphere_geometry_mod.F90:328-390): triangle whose last edge p[1]-p[2] lies on
    a latitude circle; the lune between that small circle and the great arc is added or removed."""
    res = _area(p)
    inv = lambda q: ((lambda lo: lo + _QPI2 if lo < 0 else lo)(mp.atan2(q[1], q[0])), mp.asin(q[2] / _QR))
    (lon0, lat0), (lon1, lat1), (lon2, lat2) = inv(p[0]), inv(p[1]), inv(p[2])
    assert lat1 == lat2
    if lat1 == 0:
        return res
    dlon = (lon2 - lon1) if lat0 > lat1 else (lon1 - lon2)
    if dlon < 0:
        dlon = dlon + _QPI2
    pole_side = lat0 * lat1 >= 0 and abs(lat0) > abs(lat1)
    v2, v3 = (p[1], p[2]) if pole_side else (p[2], p[1])
    if lat1 > 0:
        v1 = (_Q(0), _Q(0), _QR); area1 = _QR ** 2 * dlon * (1 - mp.sin(lat1))
    else:
        v1 = (_Q(0), _Q(0), -_QR); area1 = _QR ** 2 * dlon * (mp.sin(lat1) + 1)
    area3 = area1 - _area([v1, v2, v3])
    if area3 < 0 and abs(area3) > 1.0e-10:
        area3 = _Q(0)
    return res + area3 if pole_side else res - area3


def mesh(nlon, nlat):
    """mesh_init_global for the uniform mesh (coarse_pole_mul = 0), tangent_wgt_scheme = 'classic'.
    Returns interior rows only: full-lat vectors have nlat entries (j = 1..nlat), half-lat nlat-1."""
    nh = nlat - 1
    dlon = (PI2 - 0.0) / nlon
    full_lon = np.array([0.0 + (i - 1) * dlon for i in range(1, nlon + 1)])
    half_lon = full_lon + 0.5 * dlon
    dlat0 = (PI05 - (-PI05)) / nh
    half_lat = np.zeros(nh); full_lat = np.zeros(nlat); dlat = np.full(nh, dlat0)
    full_lat[0] = -PI05
    for j in range(1, nlat - 1):
        full_lat[j] = full_lat[j - 1] + dlat[j - 1]
        if abs(full_lat[j]) < 1.0e-12:
            full_lat[j] = 0.0
    full_lat[nlat - 1] = PI05
    for j in range(nh):
        if full_lat[j] == PI05:
            continue
        half_lat[j] = full_lat[j] + 0.5 * dlat[j]
        if abs(half_lat[j]) < 1.0e-12:
            half_lat[j] = 0.0
    for j in range(nlat):                       # equatorial symmetry (mesh_mod.F90:226-237)
        if full_lat[j] > 0:
            full_lat[j] = -full_lat[nlat - 1 - j]
    for j in range(nh):
        if half_lat[j] > 0:
            half_lat[j] = -half_lat[nh - 1 - j]
    full_cos = np.array([math.cos(x) for x in full_lat]); full_sin = np.array([math.sin(x) for x in full_lat])
    half_cos = np.array([math.cos(x) for x in half_lat]); half_sin = np.array([math.sin(x) for x in half_lat])
    full_cos[0] = 0.0; full_sin[0] = -1.0; full_cos[-1] = 0.0; full_sin[-1] = 1.0
    m = dict(dlon=dlon, full_lon=full_lon, half_lon=half_lon, full_lat=full_lat, half_lat=half_lat,
             full_lat_deg=full_lat * DEG, half_lat_deg=half_lat * DEG,
             full_cos_lat=full_cos, full_sin_lat=full_sin, half_cos_lat=half_cos, half_sin_lat=half_sin)
    z = lambda n: np.zeros(n)
    area_cell, sub = z(nlat), np.zeros((nlat, 2))
    alw, aln, als = z(nlat), z(nlat), z(nlat)
    fl = lambda i: float(full_lon[i - 1]); hl = lambda i: float(half_lon[i - 1])
    FL = lambda j: float(full_lat[j - 1]); HL = lambda j: float(half_lat[j - 1])
    for j in range(1, nlat + 1):
        if j == 1:
            area_cell[0] = RADIUS ** 2 * dlon * (half_sin[0] + 1.0)
            sub[0, 1] = RADIUS ** 2 * 0.5 * dlon * (half_sin[0] + 1.0)
        elif j == nlat:
            area_cell[j - 1] = RADIUS ** 2 * dlon * (1.0 - half_sin[j - 2])
            sub[j - 1, 0] = RADIUS ** 2 * 0.5 * dlon * (1.0 - half_sin[j - 2])
        else:
            area_cell[j - 1] = RADIUS ** 2 * dlon * (half_sin[j - 1] - half_sin[j - 2])
            sub[j - 1, 0] = RADIUS ** 2 * 0.5 * dlon * (full_sin[j - 1] - half_sin[j - 2])
            sub[j - 1, 1] = RADIUS ** 2 * 0.5 * dlon * (half_sin[j - 1] - full_sin[j - 1])
            alw[j - 1] = float(_area([_cart(fl(1), FL(j)), _cart(hl(1), HL(j - 1)), _cart(hl(1), HL(j))]))
            aln[j - 1] = float(_area_small_arc([_cart(hl(1), HL(j)), _cart(fl(1), FL(j)), _cart(fl(2), FL(j))]))
            als[j - 1] = float(_area_small_arc([_cart(hl(1), HL(j - 1)), _cart(fl(2), FL(j)), _cart(fl(1), FL(j))]))
    area_vtx, atw, atn, ats = z(nh), z(nh), z(nh), z(nh)
    for j in range(1, nh + 1):
        area_vtx[j - 1] = RADIUS ** 2 * dlon * (full_sin[j] - full_sin[j - 1])
        atw[j - 1] = float(_area([_cart(hl(1), HL(j)), _cart(fl(2), FL(j)), _cart(fl(2), FL(j + 1))]))
        atn[j - 1] = float(_area_small_arc([_cart(fl(2), FL(j + 1)), _cart(hl(1), HL(j)), _cart(hl(2), HL(j))]))
        ats[j - 1] = float(_area_small_arc([_cart(fl(2), FL(j)), _cart(hl(2), HL(j)), _cart(hl(1), HL(j))]))
        if j == 1:
            ats[j - 1] = area_cell[0]
        elif j + 1 == nlat:
            atn[j - 1] = area_cell[nlat - 1]
    area_lon = alw + alw
    area_lat = atn + ats
    de_lon, le_lon = z(nlat), z(nlat)
    for j in range(2, nlat):
        de_lon[j - 1] = RADIUS * full_cos[j - 1] * dlon
        le_lon[j - 1] = 2.0 * area_lon[j - 1] / de_lon[j - 1]
    le_lat = RADIUS * half_cos * dlon
    de_lat = 2.0 * area_lat / le_lat
    ftw, htw = np.zeros((nlat, 2)), np.zeros((nh, 2))
    for j in range(2, nlat):
        ftw[j - 1, 0] = le_lat[j - 2] / de_lon[j - 1] * 0.25
        ftw[j - 1, 1] = le_lat[j - 1] / de_lon[j - 1] * 0.25
    for j in range(1, nh + 1):
        htw[j - 1, 0] = le_lon[j - 1] / de_lat[j - 1] * 0.25
        htw[j - 1, 1] = le_lon[j] / de_lat[j - 1] * 0.25
    m.update(area_cell=area_cell, area_subcell=sub, area_lon_west=alw, area_lon_east=alw.copy(), area_lon_north=aln,
             area_lon_south=als, area_lon=area_lon, area_vtx=area_vtx, area_lat_west=atw, area_lat_east=atw.copy(),
             area_lat_north=atn, area_lat_south=ats, area_lat=area_lat, de_lon=de_lon, le_lon=le_lon, le_lat=le_lat,
             de_lat=de_lat, full_tangent_wgt=ftw, half_tangent_wgt=htw, full_f=2.0 * OMEGA * full_sin,
             half_f=2.0 * OMEGA * half_sin)
    return m


def _gaussian_weight(width, ngrid):
    s = FILTER_B * width / 2.0
    w = np.zeros(ngrid)
    for i in range(1, ngrid + 1):
        x = i - (ngrid + 1) // 2
        w[i - 1] = math.exp(-(x ** 2) / (2 * s ** 2)) / (s * math.sqrt(PI2))
    tot = 0.0
    for v in w:                                   # sum(w), left to right
        tot += v
    return w / tot


def _exp_two_values(val1, val2, x0, x1, x):
    w = math.exp((x - x0) ** 2 * math.log(1.0e-3) / (x1 - x0) ** 2)
    return w * val1 + (1 - w) * val2


def filters(m, dt_dyn, kind):
    """filter_init (filter_types_mod.F90:49-147) for kind in big_filter | small_filter1 | small_filter2.
    Returns (ngrid_lon[nlat], [wgt rows], ngrid_lat[nlat-1], [wgt rows]) on interior rows."""
    nlat = len(m["full_lat"]); nh = nlat - 1
    ngl = np.zeros(nlat, dtype=int); wdl = np.zeros(nlat); nga = np.zeros(nh, dtype=int); wda = np.zeros(nh)

    def width(dx, latdeg):
        cfl = MAX_WAVE_SPEED * dt_dyn / dx
        w = FILTER_A * cfl / MAX_CFL * (FILTER_C * (math.tanh(90 - abs(latdeg)) - 1) + 1)
        n = math.ceil(w) + 2
        if n % 2 == 0:
            n += 1
        return w, n
    for j in range(2, nlat):
        if m["de_lon"][j - 1] > 0:
            wdl[j - 1], ngl[j - 1] = width(m["de_lon"][j - 1], m["full_lat_deg"][j - 1])
    for j in range(1, nh + 1):
        if m["le_lat"][j - 1] > 0:
            wda[j - 1], nga[j - 1] = width(m["le_lat"][j - 1], m["half_lat_deg"][j - 1])
    wl = [None] * nlat; wa = [None] * nh

    def shrink(n, w, fac):
        n2 = int(fac * n)                          # integer assignment truncates
        if n2 % 2 == 0:
            n2 += 1
        return n2, fac * w
    for rows, ng, wd, out, lat0, latdeg in ((range(2, nlat), ngl, wdl, wl, -m["full_lat_deg"][1], m["full_lat_deg"]),
                                            (range(1, nh + 1), nga, wda, wa, -m["half_lat_deg"][0], m["half_lat_deg"])):
        for j in rows:
            if ng[j - 1] > 1:
                if kind == "small_filter1":
                    ng[j - 1], wd[j - 1] = shrink(ng[j - 1], wd[j - 1], FILTER_D)
                elif kind == "small_filter2":
                    fac = _exp_two_values(FILTER_E, FILTER_D, lat0, 80.0, abs(latdeg[j - 1]))
                    ng[j - 1], wd[j - 1] = shrink(ng[j - 1], wd[j - 1], fac)
                out[j - 1] = _gaussian_weight(wd[j - 1], ng[j - 1])
    return ngl, wl, nga, wa


def lon_halo_width(nga_big):
    """process_create_blocks (process_mod.F90:206-211)."""
    return max(2, max((int(n) - 1) // 2 for n in nga_big))


def vert_levels(hyai, hybi):
    """hybrid_coord_init tail (hybrid_coord_mod.F90:124-137)."""
    hyai = np.asarray(hyai, dtype=float); hybi = np.asarray(hybi, dtype=float)
    hyam = 0.5 * (hyai[:-1] + hyai[1:]); hybm = 0.5 * (hybi[:-1] + hybi[1:])
    return hyam + hybm, hyai + hybi


def ic_rossby_haurwitz(m):
    R, omg = 4.0, 7.848e-6
    gz0 = 8.0e3 * G
    nlat, nlon = len(m["full_lat"]), len(m["full_lon"])
    u = np.zeros((nlat, nlon)); v = np.zeros((nlat - 1, nlon)); gz = np.zeros((nlat, nlon))
    for j in range(nlat):
        cl, sl = m["full_cos_lat"][j], m["full_sin_lat"][j]
        for i in range(nlon):
            lon = m["half_lon"][i]
            a = cl
            b = R * cl ** (R - 1) * sl ** 2 * math.cos(R * lon)
            c = cl ** (R + 1) * math.cos(R * lon)
            u[j, i] = RADIUS * omg * (a + b - c)
    for j in range(nlat - 1):
        cl, sl = m["half_cos_lat"][j], m["half_sin_lat"][j]
        for i in range(nlon):
            lon = m["full_lon"][i]
            v[j, i] = -RADIUS * omg * (R * cl ** (R - 1) * sl * math.sin(R * lon))
    for j in range(nlat):
        cl = m["full_cos_lat"][j]
        a = 0.5 * omg * (2 * OMEGA + omg) * cl ** 2 + \
            0.25 * omg ** 2 * ((R + 1) * cl ** (2 * R + 2) + (2 * R ** 2 - R - 2) * cl ** (2 * R) - 2 * R ** 2 * cl ** (2 * R - 2))
        b = 2 * (OMEGA + omg) * omg * cl ** R * (R ** 2 + 2 * R + 2 - (R + 1) ** 2 * cl ** 2) / (R + 1) / (R + 2)
        c = 0.25 * omg ** 2 * cl ** (2 * R) * ((R + 1) * cl ** 2 - R - 2)
        for i in range(nlon):
            lon = m["full_lon"][i]
            gz[j, i] = gz0 + RADIUS ** 2 * (a + b * math.cos(R * lon) + c * math.cos(2 * R * lon))
    return dict(u=u, v=v, gz=gz, gzs=np.zeros((nlat, nlon)))


def ic_mountain_zonal_flow(m):
    u0 = f32(20.0); gz0 = 5960.0 * G; gzs0 = 2000.0 * G
    lon0 = PI * f32(1.5); lat0 = PI / f32(6.0); R = PI / f32(9.0)
    nlat, nlon = len(m["full_lat"]), len(m["full_lon"])
    gzs = np.zeros((nlat, nlon)); u = np.zeros((nlat, nlon)); v = np.zeros((nlat - 1, nlon)); gz = np.zeros((nlat, nlon))
    for j in range(nlat):
        cl, sl = m["full_cos_lat"][j], m["full_sin_lat"][j]
        for i in range(nlon):
            dlon = abs(m["full_lon"][i] - lon0)
            dlon = min(dlon, 2 * PI - dlon)
            d = min(R, math.sqrt(dlon ** 2 + (m["full_lat"][j] - lat0) ** 2))
            gzs[j, i] = gzs0 * (1.0 - d / R)
            u[j, i] = u0 * (cl * 1.0 + math.cos(m["half_lon"][i]) * sl * 0.0)
            gz[j, i] = gz0 - (RADIUS * OMEGA * u0 + u0 ** 2 * 0.5) * (sl * 1.0 - math.cos(m["full_lon"][i]) * cl * 0.0) ** 2
    return dict(u=u, v=v, gz=gz, gzs=gzs)


def ic_baroclinic_wave(m, hyai, hybi):
    """JW06 initial state (baroclinic_wave_test_mod.F90:28-170): u, pt, phs, gzs (v = 0)."""
    u0, t0, gamma, dT = 35.0, 288.0, 0.005, 4.8e5
    eta0, etat = f32(0.252), f32(0.2)
    lonc, latc, up = PI / 9.0, PI2 / 9.0, 1.0
    full_lev, half_lev = vert_levels(hyai, hybi)
    nlev = len(full_lev); nlat, nlon = len(m["full_lat"]), len(m["full_lon"])
    phs = 1.0e5
    ph_lev = np.array([hyai[k] * (P0 - PTOP) + hybi[k] * (phs - PTOP) + PTOP for k in range(nlev + 1)])
    ph = 0.5 * (ph_lev[:-1] + ph_lev[1:])
    u = np.zeros((nlev, nlat, nlon)); pt = np.zeros((nlev, nlat, nlon)); gzs = np.zeros((nlat, nlon))
    sl = m["full_sin_lat"]; cl = m["full_cos_lat"]
    pert = np.zeros((nlat, nlon))
    for j in range(nlat):
        for i in range(nlon):
            r = 10.0 * math.acos(math.sin(latc) * sl[j] + math.cos(latc) * cl[j] * math.cos(m["half_lon"][i] - lonc))
            pert[j, i] = up * (math.exp(-r ** 2) if r ** 2 <= 50 else 0.0)
    for k in range(nlev):
        eta = full_lev[k]
        etav = (eta - eta0) * PI05
        for j in range(nlat):
            u[k, j, :] = u0 * math.cos(etav) ** 1.5 * math.sin(2 * m["full_lat"][j]) ** 2 + pert[j, :]
        etav = (eta - eta0) * PI / 2.0
        if etat <= eta <= 1:
            tbar = t0 * eta ** (RD * gamma / G)
        else:
            tbar = t0 * eta ** (RD * gamma / G) + dT * (etat - eta) ** 5
        for j in range(nlat):
            t = tbar + 3.0 / 4.0 * eta * PI * u0 / RD * math.sin(etav) * math.sqrt(math.cos(etav)) * (
                (-2 * sl[j] ** 6 * (cl[j] ** 2 + 1.0 / 3.0) + 10.0 / 63.0) * 2 * u0 * math.cos(etav) ** 1.5 +
                (8.0 / 5.0 * cl[j] ** 3 * (sl[j] ** 2 + 2.0 / 3.0) - PI / 4.0) * RADIUS * OMEGA)
            pt[k, j, :] = t * (P0 / ph[k]) ** RD_O_CPD
    etav = (1 - eta0) * PI / 2
    for j in range(nlat):
        gzs[j, :] = u0 * math.cos(etav) ** 1.5 * (
            (-2 * sl[j] ** 6 * (cl[j] ** 2 + 1.0 / 3.0) + 10.0 / 63.0) * u0 * math.cos(etav) ** 1.5 +
            (8.0 / 5.0 * cl[j] ** 3 * (sl[j] ** 2 + 2.0 / 3.0) - PI / 4.0) * RADIUS * OMEGA)
    return dict(u=u, v=np.zeros((nlev, nlat - 1, nlon)), pt=pt, phs=np.full((nlat, nlon), phs), gzs=gzs)
