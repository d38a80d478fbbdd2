"""Stand-alone host for the GPU dynamics core.

Plays the part GMCORE's Fortran drivers play in production (namelist defaults,
mesh, filters, vertical coordinate, analytic initial conditions:
src/utils/namelist_mod.F90:10-122, src/dynamics/mesh_mod.F90:134-447,
src/dynamics/filter_types_mod.F90:49-147, src/tests/**) and hands the results to
the C ABI in include/gmcore_gpu.h.  The arithmetic lives in
gmcore_b200/csrc/host_*.hpp (C++ with __float128 for the spherical areas).
"""
import ctypes as C
import os

import numpy as np

from . import abi

_LIBDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")

IC_ROSSBY_HAURWITZ, IC_MOUNTAIN_ZONAL_FLOW, IC_BAROCLINIC_WAVE, IC_HELD_SUAREZ = 0, 1, 2, 3
_CASES = {"rossby_haurwitz_wave": IC_ROSSBY_HAURWITZ, "mountain_zonal_flow": IC_MOUNTAIN_ZONAL_FLOW,
          "baroclinic_wave": IC_BAROCLINIC_WAVE, "held_suarez": IC_HELD_SUAREZ}


def _load():
    path = os.path.join(_LIBDIR, "libgmcore_host.so")
    if not os.path.exists(path):
        raise RuntimeError(f"{path} missing: run `make host` (or __graft_entry__.build())")
    lib = C.CDLL(path)
    lib.gmh_create.restype = C.c_void_p
    lib.gmh_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]
    lib.gmh_create_opts.restype = C.c_void_p
    lib.gmh_create_opts.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int]
    lib.gmh_destroy.argtypes = [C.c_void_p]
    lib.gmh_hx.argtypes = [C.c_void_p]
    lib.gmh_get_vector.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]
    lib.gmh_filter_dims.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.gmh_get_filter.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double),
                                   C.POINTER(C.c_int), C.POINTER(C.c_double)]
    lib.gmh_set_ic.argtypes = [C.c_void_p, C.c_int] + [C.POINTER(C.c_double)] * 5 + [C.c_ulonglong]
    return lib


class Host:
    """Mesh + filters + namelist defaults for one run (single block owning the sphere)."""

    def __init__(self, nlon, nlat, nlev=1, dt_dyn=300.0, test_case="rossby_haurwitz_wave",
                 time_scheme="wrfrk3", ntracers=None, coarse_pole_mul=0.0, coarse_pole_decay=100.0,
                 tangent_wgt_scheme="classic", **overrides):
        self.lib = _load()
        self.nlon, self.nlat, self.nlev = nlon, nlat, nlev
        self.test_case = test_case
        ptop = 2.194e2
        if tangent_wgt_scheme not in ("classic", "thuburn09"):
            raise ValueError(f"Invalid tangent_wgt_scheme {tangent_wgt_scheme}!")
        self.h = self.lib.gmh_create_opts(nlon, nlat, nlev, float(dt_dyn), ptop, float(coarse_pole_mul),
                                          float(coarse_pole_decay), int(tangent_wgt_scheme == "thuburn09"))
        self.hx = self.lib.gmh_hx(self.h)
        self.hy = 2
        baroclinic = nlev > 1
        cfg = abi.GmcConfig()
        cfg.nlon, cfg.nlat, cfg.nlev, cfg.hx, cfg.hy = nlon, nlat, nlev, self.hx, self.hy
        cfg.lat_ibeg, cfg.lat_iend = 1, nlat
        cfg.device, cfg.rank, cfg.nranks = 0, 0, 1
        cfg.baroclinic = int(baroclinic)
        cfg.time_scheme = {"euler": abi.GMC_EULER, "pc2": abi.GMC_PC2, "wrfrk3": abi.GMC_WRFRK3}[time_scheme]
        cfg.ke_scheme = 2 if baroclinic else 1            # namelist_mod.F90:232
        cfg.pv_scheme, cfg.upwind_order_pv, cfg.pv_pole_stokes = abi.GMC_PV_UPWIND, 3, 1
        cfg.coriolis_scheme = 1
        cfg.use_div_damp = int(baroclinic)
        cfg.div_damp_order, cfg.div_damp_k0 = 2, 3
        cfg.use_smag_damp = 0
        cfg.test_forcing = abi.GMC_FORCING_HELD_SUAREZ if test_case == "held_suarez" else abi.GMC_FORCING_NONE
        cfg.ntracers = (1 if baroclinic else 0) if ntracers is None else ntracers   # moist_mod.F90:18 'qv'
        cfg.limiter_type = abi.GMC_LIMITER_MONO           # namelist_mod.F90 default 'mono'
        cfg.tvd_limiter_type = abi.GMC_TVD_VAN_LEER        # namelist_mod.F90:70 default 'van_leer'
        cfg.ffsl_flux_type = abi.GMC_FLUX_PPM              # namelist_mod.F90:69 default 'ppm'
        cfg.ke_cell_wgt, cfg.upwind_wgt_pv = 0.5, 1.0
        cfg.div_damp_coef2, cfg.div_damp_coef4 = 1.0 / 128.0, 0.01
        cfg.div_damp_top, cfg.div_damp_pole, cfg.div_damp_lat0 = 3.0, 0.0, 90.0
        cfg.smag_damp_coef = float(np.float32(0.1))
        cfg.dt_dyn = float(dt_dyn)
        cfg.dt_adv = float(dt_dyn)
        pi = float(np.arctan(1.0) * 4.0)
        cfg.omega, cfg.radius, cfg.g = 2 * pi / 86400.0, 6.37122e6, 9.80616
        cfg.Rd, cfg.cpd, cfg.p0 = 287.04, 1004.0, 1.0e5
        cfg.Rd_o_cpd = cfg.Rd / cfg.cpd
        cfg.ptop, cfg.hybrid_p0 = ptop, 1.0e5
        for k, v in overrides.items():
            if not hasattr(cfg, k):
                raise KeyError(k)
            setattr(cfg, k, v)
        self.cfg = cfg
        # mesh vectors
        self.vec = {}
        names = abi.MESH_VECTOR_NAMES + ["full_lat_deg", "half_lat_deg", "full_lon"]
        for n in names:
            ln = self.lib.gmh_get_vector(self.h, n.encode(), None)
            a = np.zeros(ln, dtype=np.float64)
            self.lib.gmh_get_vector(self.h, n.encode(), abi.dptr(a))
            self.vec[n] = a
        mv = abi.GmcMeshVectors()
        for n in abi.MESH_VECTOR_NAMES:
            lev = n in ("hyai", "hybi", "full_lev", "half_lev", "full_dlev", "half_dlev")
            setattr(mv, n, abi.dptr(None if (lev and not baroclinic) else self.vec[n]))
        self.mesh_vectors = mv
        # filters
        self.filters = []
        nf, nh = nlat + 2 * self.hy, nlat - 1 + 2 * self.hy
        for which in range(3):
            ldl, lda = C.c_int(), C.c_int()
            self.lib.gmh_filter_dims(self.h, which, C.byref(ldl), C.byref(lda))
            ngl = np.zeros(nf, dtype=np.int32); nga = np.zeros(nh, dtype=np.int32)
            wl = np.zeros(nf * ldl.value, dtype=np.float64); wa = np.zeros(nh * lda.value, dtype=np.float64)
            self.lib.gmh_get_filter(self.h, which, abi.iptr(ngl), abi.dptr(wl), abi.iptr(nga), abi.dptr(wa))
            self.filters.append((ngl, wl, ldl.value, nga, wa, lda.value))

    # ---- halo'd Fortran storage shapes (C order = (lev, lat, lon)) ----
    @property
    def nx(self):
        return self.nlon + 2 * self.hx

    def shape(self, loc):
        """loc: 'cell'/'lon' (full lat), 'lat'/'vtx' (half lat), '+lev' suffix for half levels, '2d'."""
        nyf, nyh = self.nlat + 2 * self.hy, self.nlat - 1 + 2 * self.hy
        if loc == "2d":
            return (nyf, self.nx)
        ny = nyf if loc.split("+")[0] in ("cell", "lon") else nyh
        nz = self.nlev + 1 if loc.endswith("+lev") else self.nlev + 4
        return (nz, ny, self.nx)

    def interior(self, a, loc):
        """View of the interior of a halo'd array."""
        hx, hy = self.hx, self.hy
        if loc == "2d":
            return a[hy:hy + self.nlat, hx:hx + self.nlon]
        nj = self.nlat if loc.split("+")[0] in ("cell", "lon") else self.nlat - 1
        if loc.endswith("+lev"):
            return a[:, hy:hy + nj, hx:hx + self.nlon]
        return a[2:2 + self.nlev, hy:hy + nj, hx:hx + self.nlon]

    def initial_condition(self, seed=20261104):
        u = np.zeros(self.shape("lon")); v = np.zeros(self.shape("lat"))
        s = np.zeros(self.shape("cell")); phs = np.zeros(self.shape("2d")); gzs = np.zeros(self.shape("2d"))
        rc = self.lib.gmh_set_ic(self.h, _CASES[self.test_case], abi.dptr(u), abi.dptr(v), abi.dptr(s),
                                 abi.dptr(phs), abi.dptr(gzs), seed)
        if rc != 0:
            raise RuntimeError(f"unknown test case {self.test_case}")
        return {"u_lon": u, "v_lat": v, "pt_or_gz": s, "phs": phs if self.nlev > 1 else None, "gzs": gzs}

    def __del__(self):
        try:
            self.lib.gmh_destroy(self.h)
        except Exception:
            pass
