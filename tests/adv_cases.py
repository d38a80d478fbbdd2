"""The reference's stand-alone advection cases beyond solid rotation, as prescribed-wind generators
for the advection-only driver loop (drivers/gmcore_adv_driver.F90:84-88: set_uv(new, t + dt) -> adv_run(new)):

  deform_case4     Nair & Lauritzen deformational flow + solid-body rotation, four tracers (background, cosine hills,
                   slotted cylinders, Gaussian hills); back at the start after one period of 12 days
                   (src/tests/adv/deform_test_mod.F90:28-142,248-289)
  moving_vortices  two vortices carried around the sphere over the poles, with the analytic solution
                   (src/tests/adv/moving_vortices_test_mod.F90:36-147; Nair & Jablonowski 2008)
  dcmip12          DCMIP 1-2: 3-D Hadley-like circulation reversing at half time, prescribed vertical mass flux we_lev
                   (src/tests/adv/dcmip12_test_mod.F90:20-161)

The winds are formed here in numpy from the same formulas; the oracle and the CUDA path receive identical arrays
(gmc_adv_prescribe / gmc_adv_prescribe_we), so GPU-vs-oracle parity is per step, and the oracle is held to the
analytic properties of each case in tests/test_advection_cpu.py.
"""
import numpy as np

from gmcore_b200.host import Host

DAY = 86400.0
PI = np.pi


def _grids(host):
    hx, hy, nlon, nlat = host.hx, host.hy, host.nlon, host.nlat
    return (host.vec["full_lon"][hx:hx + nlon], host.vec["half_lon"][hx:hx + nlon],
            host.vec["full_lat"][hy:hy + nlat], host.vec["half_lat"][hy:hy + nlat - 1])


def _put(host, arr, loc, val):
    host.interior(arr, loc)[...] = val
    return arr


def _dist(lon1, lat1, lon, lat):      # calc_distance / radius (sphere_geometry_mod.F90)
    return np.arccos(np.clip(np.sin(lat1) * np.sin(lat) + np.cos(lat1) * np.cos(lat) * np.cos(lon1 - lon), -1.0, 1.0))


def _rot(lon_p, lat_p, lon_o, lat_o):      # rotation_transform (sphere_geometry_mod.F90:168-191)
    dlon = lon_o - lon_p
    lon_r = np.arctan2(np.cos(lat_o) * np.sin(dlon), np.cos(lat_o) * np.sin(lat_p) * np.cos(dlon) - np.cos(lat_p) * np.sin(lat_o))
    lon_r = np.where(lon_r < 0, 2 * PI + lon_r, lon_r)
    lat_r = np.arcsin(np.clip(np.sin(lat_o) * np.sin(lat_p) + np.cos(lat_o) * np.cos(lat_p) * np.cos(dlon), -1, 1))
    return lon_r, lat_r


def _inv_rot(lon_p, lat_p, lon_r, lat_r):  # inverse_rotation_transform (:218-247)
    t1 = np.cos(lat_r) * np.sin(lon_r)
    t2 = np.sin(lat_r) * np.cos(lat_p) + np.cos(lat_r) * np.cos(lon_r) * np.sin(lat_p)
    if abs(t2) < np.finfo(float).eps:
        t2 = 0.0
    lon_o = lon_p + np.arctan2(t1, t2)
    if lon_o > 2 * PI:
        lon_o -= 2 * PI
    lat_o = np.arcsin(np.clip(np.sin(lat_r) * np.sin(lat_p) - np.cos(lat_r) * np.cos(lat_p) * np.cos(lon_r), -1, 1))
    return lon_o, lat_o


class DeformCase4:
    period = 12 * DAY

    def __init__(self, nlon=90, nlat=46, nlev=1, dt=3600.0):
        self.host = Host(nlon, nlat, max(nlev, 2), dt, "baroclinic_wave", ntracers=4, use_div_damp=0)
        self.dt = dt
        self.nsteps = int(round(self.period / dt))

    def tracers(self):
        host = self.host; a = host.cfg.radius
        fl, hl, fa, ha = _grids(host)
        lat, lon = np.meshgrid(fa, fl, indexing="ij")
        lon1, lat1, lon2, lat2 = PI * 5 / 6, 0.0, PI * 7 / 6, 0.0
        q = np.zeros((4,) + host.shape("cell"))
        _put(host, q[0], "cell", 1.0)
        r = 0.5                                            # r / radius
        r1, r2 = _dist(lon1, lat1, lon, lat), _dist(lon2, lat2, lon, lat)
        hills = np.where(r1 < r, 0.1 + 0.9 * 0.5 * (1 + np.cos(PI * r1 / r)), np.where(r2 < r, 0.1 + 0.9 * 0.5 * (1 + np.cos(PI * r2 / r)), 0.1))
        _put(host, q[1], "cell", hills[None])
        cyl = np.full_like(lat, 0.1)
        c1 = ((r1 <= r) & (np.abs(lon - lon1) >= r / 6)) | ((r2 <= r) & (np.abs(lon - lon2) >= r / 6))
        c2 = (r1 <= r) & (np.abs(lon - lon1) < r / 6) & (lat - lat1 < -5.0 / 12.0 * r)
        c3 = (r2 <= r) & (np.abs(lon - lon2) < r / 6) & (lat - lat2 > 5.0 / 12.0 * r)
        cyl[c1 | c2 | c3] = 1.0
        _put(host, q[2], "cell", cyl[None])
        xyz = lambda lo, la: np.stack([np.cos(la) * np.cos(lo), np.cos(la) * np.sin(lo), np.sin(la)])
        x = xyz(lon, lat); x1 = xyz(lon1, lat1)[:, None, None]; x2 = xyz(lon2, lat2)[:, None, None]
        gauss = 0.95 * (np.exp(-5.0 * ((x - x1) ** 2).sum(0)) + np.exp(-5.0 * ((x - x2) ** 2).sum(0)))
        _put(host, q[3], "cell", gauss[None])
        self.q0 = [host.interior(q[l], "cell")[0].copy() for l in range(4)]
        return q

    def winds(self, t):
        host = self.host; a = host.cfg.radius
        fl, hl, fa, ha = _grids(host)
        k = 10.0 * a / self.period; c1 = 2 * PI * t / self.period; c2 = 2 * PI * a / self.period
        cos_t = np.cos(PI * t / self.period)
        lat, lon = np.meshgrid(fa, hl - c1, indexing="ij")
        uu = k * np.sin(lon) ** 2 * np.sin(2 * lat) * cos_t + c2 * np.cos(lat)
        uu[0] = 0; uu[-1] = 0
        lat, lon = np.meshgrid(ha, fl - c1, indexing="ij")
        vv = k * np.sin(2 * lon) * np.cos(lat) * cos_t
        u = _put(host, np.zeros(host.shape("lon")), "lon", uu[None]); v = _put(host, np.zeros(host.shape("lat")), "lat", vv[None])
        return u, v, None


class MovingVortices:
    period = 12 * DAY
    rho0, gamma, alpha = 3.0, 5.0, PI / 2
    lonp0, latp0, lonv0, latv0 = PI, 0.0, 1.5 * PI, 0.0          # latp0 = pi/2 - alpha

    def __init__(self, nlon=90, nlat=46, nlev=1, dt=3600.0):
        self.host = Host(nlon, nlat, max(nlev, 2), dt, "baroclinic_wave", ntracers=2, use_div_damp=0)
        self.dt = dt
        self.nsteps = int(round(self.period / dt))
        self.u0 = 2 * PI * self.host.cfg.radius / self.period
        self.lonvr0, _ = _rot(self.lonp0, self.latp0, self.lonv0, self.latv0)

    def _a_omg(self, latr):
        r = self.rho0 * np.cos(latr)
        c = 1.5 * np.sqrt(3.0)
        safe = np.where(np.abs(r) < np.finfo(float).eps, 1.0, r)
        return np.where(np.abs(r) < np.finfo(float).eps, 0.0, self.u0 * c / np.cosh(safe) ** 2 * np.tanh(safe) / safe)

    def vortex_centre(self, t):
        lonvr = self.lonvr0 + self.u0 / self.host.cfg.radius * t
        if lonvr > 2 * PI:
            lonvr -= 2 * PI
        _, latvr = _rot(self.lonp0, self.latp0, self.lonv0, self.latv0)
        return _inv_rot(self.lonp0, self.latp0, lonvr, latvr)

    def tracers(self):
        host = self.host
        fl, hl, fa, ha = _grids(host)
        q = np.zeros((2,) + host.shape("cell"))
        _put(host, q[0], "cell", 1.0)
        _put(host, q[1], "cell", self.analytic(0.0)[None])
        return q

    def analytic(self, t):
        """q1(t): the vortex field around the moved centre with the accumulated rotation (Nair & Jablonowski 2008, eq. 23)."""
        host = self.host
        fl, hl, fa, ha = _grids(host)
        lat, lon = np.meshgrid(fa, fl, indexing="ij")
        lonv, latv = self.vortex_centre(t)
        lonr, latr = _rot(lonv, latv, lon, lat)
        r = self.rho0 * np.cos(latr)
        omg = self._a_omg(latr) / host.cfg.radius
        return 1 - np.tanh(r / self.gamma * np.sin(lonr - omg * t))

    def winds(self, t):
        host = self.host
        fl, hl, fa, ha = _grids(host)
        lonv, latv = self.vortex_centre(t)
        lat, lon = np.meshgrid(fa, hl, indexing="ij")
        _, latr = _rot(lonv, latv, lon, lat)
        uu = self.u0 * (np.cos(lat) * np.cos(self.alpha) + np.sin(lat) * np.cos(lon) * np.sin(self.alpha)) + \
            self._a_omg(latr) * (np.sin(latv) * np.cos(lat) - np.cos(latv) * np.cos(lon - lonv) * np.sin(lat))
        uu[0] = 0; uu[-1] = 0
        lat, lon = np.meshgrid(ha, fl, indexing="ij")
        _, latr = _rot(lonv, latv, lon, lat)
        vv = -self.u0 * np.sin(lon) * np.sin(self.alpha) + self._a_omg(latr) * np.cos(latv) * np.sin(lon - lonv)
        u = _put(host, np.zeros(host.shape("lon")), "lon", uu[None]); v = _put(host, np.zeros(host.shape("lat")), "lat", vv[None])
        return u, v, None


class Dcmip12:
    T0, p0, K0, u0, w0, z1, z2, ztop, tau = 300.0, 1.0e5, 5.0, 40.0, 0.15, 2000.0, 5000.0, 12000.0, DAY

    def __init__(self, nlon=90, nlat=46, nlev=16, dt=1800.0):
        # equal_eta levels with the case's model top (dcmip12_test_init: ptop = 25494.4)
        self.host = Host(nlon, nlat, nlev, dt, "baroclinic_wave", ntracers=2, use_div_damp=0, ptop=25494.4)
        self.dt = dt
        self.nsteps = int(round(self.tau / dt))
        host = self.host; c = host.cfg
        hyai, hybi = host.vec["hyai"], host.vec["hybi"]
        self.ph_lev = hyai * (c.hybrid_p0 - c.ptop) + hybi * (self.p0 - c.ptop) + c.ptop            # phs = p0 everywhere
        self.ph = 0.5 * (self.ph_lev[:-1] + self.ph_lev[1:])
        # calc_gz_lev with t = T0, gzs = 0: the sum of Rd T0 log ratios telescopes
        self.gz_lev = np.array([sum(c.Rd * self.T0 * np.log(self.ph_lev[l + 1] / self.ph_lev[l]) for l in range(k, nlev)) for k in range(nlev + 1)])
        self.gz = 0.5 * (self.gz_lev[:-1] + self.gz_lev[1:])
        self.rho0 = self.p0 / (c.Rd * self.T0)

    def tracers(self):
        host = self.host
        q = np.zeros((2,) + host.shape("cell"))
        _put(host, q[0], "cell", 1.0)
        z = self.gz / host.cfg.g
        z0 = 0.5 * (self.z1 + self.z2)
        prof = np.where((self.z1 < z) & (z < self.z2), 0.5 * (1 + np.cos(2 * PI * (z - z0) / (self.z2 - self.z1))), 0.0)
        _put(host, q[1], "cell", prof[:, None, None] * np.ones((1, host.nlat, host.nlon)))
        self.q1_0 = host.interior(q[1], "cell").copy()
        return q

    def winds(self, t):
        host = self.host; c = host.cfg; a = c.radius; nlev = host.nlev
        fl, hl, fa, ha = _grids(host)
        cos_t = np.cos(PI * t / self.tau)
        uu = self.u0 * np.cos(fa)[None, :, None] * np.ones((nlev, 1, host.nlon))
        uu[:, 0] = 0; uu[:, -1] = 0
        rho = self.ph / (c.Rd * self.T0)
        vv = (-a * self.w0 * PI * self.rho0 / (self.K0 * self.ztop * rho))[:, None, None] * \
            (np.cos(ha) * np.sin(self.K0 * ha))[None, :, None] * np.cos(PI * self.gz / (c.g * self.ztop))[:, None, None] * cos_t * np.ones((1, 1, host.nlon))
        half_dlev = host.vec["half_dlev"]
        we = np.zeros(host.shape("cell+lev"))
        rho_l = self.ph_lev / (c.Rd * self.T0)
        lat = fa
        for k in range(1, nlev):                      # half levels 2 .. nlev (m_lev = 1)
            dphdlev = 1.0 / half_dlev[k]
            w = -dphdlev * c.g * self.w0 * self.rho0 / self.K0 * (-2 * np.sin(self.K0 * lat) * np.sin(lat) + self.K0 * np.cos(lat) * np.cos(self.K0 * lat)) * \
                np.sin(PI * self.gz_lev[k] / (c.g * self.ztop)) * cos_t / self.p0
            host.interior(we, "cell+lev")[k] = w[:, None]
        u = _put(host, np.zeros(host.shape("lon")), "lon", uu); v = _put(host, np.zeros(host.shape("lat")), "lat", vv)
        return u, v, we


def run_case(model, case, nsteps, q):
    """gmcore_adv_driver loop: set_uv(new, t + dt) -> adv_run(new) -> time_advance.  adv_run does not advect while
    time_step == 0 (adv_mod.F90:166), hence one extra call, as in the solid-rotation runner."""
    model.set_tracers(q)
    for s in range(nsteps + 1):
        u, v, we = case.winds((s + 1) * case.dt)
        model.adv_prescribe(u, v, we_lev=we)
        model.adv_step()
    return model.get_tracers()
