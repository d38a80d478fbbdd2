"""Python mirror of the GMCORE driver interface on top of the C ABI.

`GmcoreGPU` plays gmcore_init / gmcore_run (src/dynamics/gmcore_mod.F90:45-149)
for a stand-alone run: every call goes through libgmcore_b200.so (ctypes, plain
pointers); there is no CPU path — constructing it without a usable CUDA device
raises.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_LIBDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
_LIB = None

ABI_SYMBOLS = ["gmc_create", "gmc_destroy", "gmc_set_mesh", "gmc_set_filters", "gmc_set_static",
               "gmc_upload_state", "gmc_download_state", "gmc_set_tracers", "gmc_get_tracers",
               "gmc_prepare_run", "gmc_time_integrate", "gmc_run_steps", "gmc_diagnose", "gmc_get_field",
               "gmc_synchronize", "gmc_launch_count", "gmc_last_error", "gmc_reset_clock", "gmc_adv_prescribe", "gmc_adv_prescribe_we", "gmc_adv_step", "gmc_interp_pressure_level", "gmc_diagnose_enqueue", "gmc_diagnose_collect", "gmc_diagnose_static", "gmc_fastmath_eval",
               "gmc_group_link", "gmc_nccl_unique_id", "gmc_nccl_init", "gmc_pin_host", "gmc_unpin_host",
               "gmc_add_tracer_batch", "gmc_set_batch_tracers", "gmc_get_batch_tracers"]


def load_library():
    global _LIB
    if _LIB is not None:
        return _LIB
    # GMCORE_B200_LIB: alternative build of the same library (kernel A/B experiments, tools/kprof.py)
    path = os.environ.get("GMCORE_B200_LIB") or os.path.join(_LIBDIR, "libgmcore_b200.so")
    if not os.path.exists(path):
        raise RuntimeError(f"{path} missing: the CUDA extension is not built (make gpu); no CPU fallback exists")
    lib = C.CDLL(path)
    P, I = C.POINTER(C.c_double), C.POINTER(C.c_int)
    V = C.c_void_p
    lib.gmc_create.argtypes = [C.POINTER(abi.GmcConfig), C.POINTER(V)]
    lib.gmc_create_error.restype = C.c_char_p
    lib.gmc_destroy.argtypes = [V]
    lib.gmc_set_mesh.argtypes = [V, C.POINTER(abi.GmcMeshVectors)]
    lib.gmc_set_filters.argtypes = [V, C.c_int, I, P, C.c_int, I, P, C.c_int]
    lib.gmc_set_static.argtypes = [V, P]
    lib.gmc_upload_state.argtypes = [V, C.c_int, P, P, P, P]
    lib.gmc_download_state.argtypes = [V, C.c_int, P, P, P, P]
    lib.gmc_set_tracers.argtypes = [V, P]
    lib.gmc_get_tracers.argtypes = [V, P]
    lib.gmc_add_tracer_batch.argtypes = [V, C.c_int, C.c_double, I]
    lib.gmc_set_batch_tracers.argtypes = [V, C.c_int, P]
    lib.gmc_get_batch_tracers.argtypes = [V, C.c_int, P]
    lib.gmc_prepare_run.argtypes = [V, C.c_int]
    lib.gmc_time_integrate.argtypes = [V, C.c_int, C.c_int, C.c_double, C.c_int]
    lib.gmc_run_steps.argtypes = [V, I, I, I, C.c_int, P]
    lib.gmc_diagnose.argtypes = [V, C.c_int, P]
    lib.gmc_get_field.argtypes = [V, C.c_int, C.c_char_p, P]
    lib.gmc_synchronize.argtypes = [V]
    lib.gmc_launch_count.argtypes = [V, C.POINTER(C.c_longlong)]
    lib.gmc_last_error.argtypes = [V, C.c_char_p, C.c_int]
    lib.gmc_struct_sizes.argtypes = [I, I]
    lib.gmc_group_link.argtypes = [C.POINTER(V), C.c_int]
    lib.gmc_nccl_unique_id.argtypes = [C.c_char_p]
    lib.gmc_nccl_init.argtypes = [V, C.c_char_p]
    lib.gmc_timer_record.argtypes = [V, C.c_int]
    lib.gmc_timer_elapsed_ms.argtypes = [V, P]
    lib.gmc_profile.argtypes = [V, C.c_int]
    lib.gmc_profile_report.argtypes = [V, C.c_char_p, C.c_int]
    lib.gmc_reset_clock.argtypes = [V]
    lib.gmc_pin_host.argtypes = [V, C.c_void_p, C.c_size_t]
    lib.gmc_unpin_host.argtypes = [V, C.c_void_p]
    lib.gmc_adv_prescribe.argtypes = [V, C.c_int, P, P]
    lib.gmc_adv_prescribe_we.argtypes = [V, C.c_int, P]
    lib.gmc_diagnose_enqueue.argtypes = [V, C.c_int]
    lib.gmc_diagnose_static.argtypes = [V, C.c_int, P]
    lib.gmc_fastmath_eval.argtypes = [C.c_int, P, C.c_double, P, C.c_int]
    lib.gmc_diagnose_collect.argtypes = [V, P, C.c_int, C.POINTER(C.c_int)]
    lib.gmc_interp_pressure_level.argtypes = [V, C.c_int, C.c_char_p, C.c_double, C.c_int, C.c_double, P]
    lib.gmc_adv_step.argtypes = [V, C.c_int, C.c_int]
    lib.gmc_stream.restype = C.c_void_p
    lib.gmc_stream.argtypes = [V]
    _LIB = lib
    return lib


# grid location of the named fields (for shapes of gmc_get_field buffers)
FIELD_LOC = {
    "u_lon": "lon", "u_f": "lon", "v_lat": "lat", "v_f": "lat", "gz": "cell", "gz_f": "cell", "m": "cell",
    "mfx_lon": "lon", "mfy_lat": "lat", "u_lat": "lat", "v_lon": "lon", "ke": "cell", "pv": "vtx",
    "pv_lon": "lon", "pv_lat": "lat", "div": "cell", "pt": "cell", "t": "cell", "phs": "2d", "phs_f": "2d",
    "ph_lev": "cell+lev", "ph_exn_lev": "cell+lev", "gz_lev": "cell+lev", "we_lev": "cell+lev", "m_lev": "cell+lev",
    "du": "lon", "dv": "lat", "gzs": "2d", "cflx": "lon", "cfly": "lat", "cflz": "cell+lev", "divx": "cell",
    "divy": "cell", "qx": "cell", "qy": "cell", "ptf_lon": "lon", "ptf_lat": "lat", "dpt": "cell",
    "wedudlev": "lon", "wedvdlev": "lat", "dphs": "2d", "ph": "cell",
}


class GmcoreGPU:
    def __init__(self, host, device=0):
        self.lib = load_library()
        self.host = host
        host.cfg.device = device
        self.h = C.c_void_p()
        rc = self.lib.gmc_create(C.byref(host.cfg), C.byref(self.h))
        if rc != 0:
            raise RuntimeError("gmc_create failed: " + self.lib.gmc_create_error().decode())
        self._chk(self.lib.gmc_set_mesh(self.h, C.byref(host.mesh_vectors)))
        for which, (ngl, wl, ldl, nga, wa, lda) in enumerate(host.filters):
            self._chk(self.lib.gmc_set_filters(self.h, which, abi.iptr(ngl), abi.dptr(wl), ldl, abi.iptr(nga), abi.dptr(wa), lda))
        self.old, self.new, self.time_step = C.c_int(1), C.c_int(2), C.c_int(0)

    def _chk(self, rc):
        if rc != 0:
            buf = C.create_string_buffer(512)
            self.lib.gmc_last_error(self.h, buf, 512)
            raise RuntimeError("gmcore_b200: " + buf.value.decode())

    # gmcore_run pre-loop ------------------------------------------------------------
    def set_state(self, ic, itime=1):
        self._chk(self.lib.gmc_set_static(self.h, abi.dptr(ic["gzs"])))
        self._chk(self.lib.gmc_upload_state(self.h, itime, abi.dptr(ic["u_lon"]), abi.dptr(ic["v_lat"]),
                                            abi.dptr(ic["pt_or_gz"]), abi.dptr(ic.get("phs"))))

    def add_tracer_batch(self, ntracers, dt_adv):
        """adv_add_tracer + adv_allocate_tracers (adv_mod.F90:314-418): one more batch; returns its index."""
        b = C.c_int(-1)
        self._chk(self.lib.gmc_add_tracer_batch(self.h, ntracers, float(dt_adv), C.byref(b)))
        self._batch_ntr()[b.value] = ntracers
        return b.value

    def _batch_ntr(self):
        if not hasattr(self, "batch_ntr"):
            n = self.host.cfg.ntracers
            self.batch_ntr = {0: n} if n > 0 else {}
        return self.batch_ntr

    def set_tracers(self, q, batch=0):
        assert q.shape[0] == self._batch_ntr()[batch]
        self._chk(self.lib.gmc_set_batch_tracers(self.h, batch, abi.dptr(q)))

    def get_tracers(self, batch=0):
        q = np.zeros((self._batch_ntr()[batch],) + self.host.shape("cell"))
        self._chk(self.lib.gmc_get_batch_tracers(self.h, batch, abi.dptr(q)))
        return q

    # advection-only driver (gmcore_adv_driver.F90:84-88): set_uv(new) -> adv_run(new) -> time_advance
    def adv_prescribe(self, u_f, v_f, itime=None, we_lev=None):
        it = self.new.value if itime is None else itime
        self._chk(self.lib.gmc_adv_prescribe(self.h, it, abi.dptr(u_f), abi.dptr(v_f)))
        if we_lev is not None:
            self._chk(self.lib.gmc_adv_prescribe_we(self.h, it, abi.dptr(we_lev)))

    def adv_step(self):
        self._chk(self.lib.gmc_adv_step(self.h, self.new.value, self.time_step.value))
        self.old.value, self.new.value = self.new.value, self.old.value
        self.time_step.value += 1

    def reset_clock(self):
        """time_init: step counter 0, old/new = 1/2, tracer-batch sample counters cleared (before a fresh upload)."""
        self._chk(self.lib.gmc_reset_clock(self.h))
        self.old.value, self.new.value, self.time_step.value = 1, 2, 0

    def prepare_run(self):
        self._chk(self.lib.gmc_prepare_run(self.h, self.old.value))

    # time loop ------------------------------------------------------------------------
    def run_steps(self, n, diag=True):
        d = np.zeros((n, 8)) if diag else None
        self._chk(self.lib.gmc_run_steps(self.h, C.byref(self.old), C.byref(self.new), C.byref(self.time_step), n,
                                         abi.dptr(d) if diag else None))
        return d

    def time_integrate(self):
        """One reference-style iteration: time_integrate + time_advance (host-side swap)."""
        self._chk(self.lib.gmc_time_integrate(self.h, self.old.value, self.new.value, self.host.cfg.dt_dyn, self.time_step.value))
        self.old.value, self.new.value = self.new.value, self.old.value
        self.time_step.value += 1

    def diagnose(self, itime=None):
        out = np.zeros(8)
        self._chk(self.lib.gmc_diagnose(self.h, self.old.value if itime is None else itime, abi.dptr(out)))
        return out

    def diagnose_static(self, itime=None):
        """gmc_diagnose_static: the same sums with the scratch in the library's static device array (no allocation)."""
        out = np.zeros(8)
        self._chk(self.lib.gmc_diagnose_static(self.h, self.old.value if itime is None else itime, abi.dptr(out)))
        return out

    def diagnose_enqueue(self, itime=None):
        """diagnose of state(itime) into the device ring: no host round trip (gmc_diagnose_collect fetches them)."""
        self._chk(self.lib.gmc_diagnose_enqueue(self.h, self.old.value if itime is None else itime))

    def diagnose_collect(self, max_sets=4096):
        out = np.zeros((max_sets, 8)); n = C.c_int(0)
        self._chk(self.lib.gmc_diagnose_collect(self.h, abi.dptr(out), max_sets, C.byref(n)))
        return out[:n.value].copy()

    def interp_pressure_level(self, var, po, logp=False, fill=np.nan, itime=None):
        """diag_state's pressure-level interpolation of one variable ("u", "v", "pt", "t", "z") -> halo'd 2-D array."""
        y = np.zeros(self.host.shape("2d"))
        self._chk(self.lib.gmc_interp_pressure_level(self.h, self.old.value if itime is None else itime, var.encode(),
                                                     float(po), int(logp), float(fill), abi.dptr(y)))
        return y

    def synchronize(self):
        self._chk(self.lib.gmc_synchronize(self.h))

    @property
    def launches(self):
        n = C.c_longlong()
        self.lib.gmc_launch_count(self.h, C.byref(n))
        return n.value

    def timer_start(self):
        self._chk(self.lib.gmc_timer_record(self.h, 0))

    def timer_stop_ms(self):
        self._chk(self.lib.gmc_timer_record(self.h, 1))
        ms = C.c_double()
        self._chk(self.lib.gmc_timer_elapsed_ms(self.h, C.byref(ms)))
        return ms.value

    def profile(self, on):
        self._chk(self.lib.gmc_profile(self.h, int(on)))

    def profile_report(self):
        """{kernel: (launches, total_ms)} for the launches since profile(True)."""
        buf = C.create_string_buffer(1 << 16)
        self._chk(self.lib.gmc_profile_report(self.h, buf, len(buf)))
        out = {}
        for line in buf.value.decode().splitlines():
            n, c, ms = line.split()
            out[n] = (int(c), float(ms))
        return out

    @property
    def stream(self):
        return self.lib.gmc_stream(self.h)

    # read-back --------------------------------------------------------------------------
    def pin_host(self, a):
        """Page-lock a host array used for repeated uploads / downloads (gmc_pin_host)."""
        self._chk(self.lib.gmc_pin_host(self.h, a.ctypes.data, a.nbytes))

    def unpin_host(self, a):
        self._chk(self.lib.gmc_unpin_host(self.h, a.ctypes.data))

    def download_state(self, itime=None, out=None):
        """out: optional dict of preallocated halo'd host arrays (u_lon, v_lat, pt_or_gz, phs) -- e.g. the host
        model's own page-locked state arrays -- written in place instead of fresh pageable ones."""
        itime = self.old.value if itime is None else itime
        hst = self.host
        if out is not None:
            u, v, s, phs = out["u_lon"], out["v_lat"], out["pt_or_gz"], out.get("phs")
            for a, loc in ((u, "lon"), (v, "lat"), (s, "cell")) + (((phs, "2d"),) if phs is not None else ()):
                if a.shape != tuple(hst.shape(loc)) or a.dtype != np.float64 or not a.flags.c_contiguous:
                    raise ValueError(f"download_state: out array for '{loc}' must be C-contiguous float64 of shape {tuple(hst.shape(loc))}")
        else:
            u = np.zeros(hst.shape("lon")); v = np.zeros(hst.shape("lat")); s = np.zeros(hst.shape("cell"))
            phs = np.zeros(hst.shape("2d")) if hst.nlev > 1 else None
        self._chk(self.lib.gmc_download_state(self.h, itime, abi.dptr(u), abi.dptr(v), abi.dptr(s), abi.dptr(phs)))
        return {"u_lon": u, "v_lat": v, "pt_or_gz": s, "phs": phs}

    def field(self, name, itime=None):
        itime = self.old.value if itime is None else itime
        a = np.zeros(self.host.shape(FIELD_LOC[name]))
        self._chk(self.lib.gmc_get_field(self.h, itime, name.encode(), abi.dptr(a)))
        return a

    def close(self):
        if self.h:
            self.lib.gmc_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- latitude bands (1-D latitude decomposition, process_mod.F90:93-139) -------------------
def round_robin(n, nparts):
    """Rows per part as parallel_types_mod.F90:235-262: the remainder goes to the lowest ranks.
    Returns [(ibeg, iend)] 1-based inclusive."""
    base, rem = divmod(n, nparts)
    out, beg = [], 1
    for r in range(nparts):
        cnt = base + (1 if r < rem else 0)
        out.append((beg, beg + cnt - 1)); beg += cnt
    return out


# Cost model of one latitude row for the optional balanced split: 1 (everything uniform in latitude) + ROW_TAP_COST
# per zonal filter tap above the equatorial count (big_filter on u, v; small_filter2 on u, v; small_filter1 on pt) +
# POLE_ROW_COST for a pole row (zonal sums, k_pole1 / k_pole_sum launches).  The weights came from rank 0's kernel
# times of the 8-band 0.25 deg x L32 run (+0.36 ms in k_filter_rows, +0.17 ms in pole kernels).  MEASURED: the split
# they give (82/92/93/93/93/93/92/82) ran at 7497 SDPD against 7565 for 8 x 90 rows -- rank 0 got 6 % faster but the
# interior bands, now 105 rows with their halo, became the critical path: at ~100 rows per GPU the kernels no longer
# fill the device and the cost per row is far from uniform, so the modelled polar surplus was ~5x too large.
# Equal rows stay the default; the option is kept for hosts with much wider polar filters (finer grids).
ROW_TAP_COST = 0.00486
POLE_ROW_COST = 5.0
BOUNDARY_COST = 4.0      # row-equivalents of redundant halo-row compute per interior band boundary (6 rows, not every kernel)


def row_weights(host):
    big_l, _, _, big_a, _, _ = host.filters[0]
    s1_l = host.filters[1][0]
    s2_l, _, _, s2_a, _, _ = host.filters[2]
    hy = host.hy
    T = np.zeros(host.nlat)
    for j in range(1, host.nlat + 1):
        t = int(big_l[j - 1 + hy]) + int(s1_l[j - 1 + hy]) + int(s2_l[j - 1 + hy])
        if j <= host.nlat - 1:
            t += int(big_a[j - 1 + hy]) + int(s2_a[j - 1 + hy])
        T[j - 1] = t
    w = 1.0 + ROW_TAP_COST * np.maximum(T - np.median(T), 0.0)
    w[0] += POLE_ROW_COST; w[-1] += POLE_ROW_COST
    return w


def balanced_rows(host, nparts, min_rows=6):
    """Contiguous latitude bands of equal modelled cost (the polar bands get fewer rows: their rows carry the wide
    zonal filters).  The reference's rule is round_robin (equal row counts); the band boundaries are the host's
    choice -- gmc_create takes lat_ibeg / lat_iend -- and results do not depend on them (tools/band_check.py).
    Returns [(ibeg, iend)] 1-based inclusive."""
    n = host.nlat
    if nparts == 1:
        return [(1, n)]
    cw = np.concatenate([[0.0], np.cumsum(row_weights(host))])
    target = (cw[-1] + BOUNDARY_COST * 2 * (nparts - 1)) / nparts
    cuts = [0]
    for k in range(1, nparts):
        want = k * target - BOUNDARY_COST * (2 * k - 1)          # rows of bands 0..k-1 (which have 2k-1 boundaries)
        j = int(np.searchsorted(cw, want))                       # first prefix >= want
        if j > 0 and want - cw[j - 1] < cw[j] - want:
            j -= 1                                               # nearest prefix
        j = max(j, cuts[-1] + min_rows)
        j = min(j, n - min_rows * (nparts - k))
        cuts.append(j)
    cuts.append(n)
    return [(cuts[k] + 1, cuts[k + 1]) for k in range(nparts)]


def band_config(host, rank, nranks, device, rows="round_robin"):
    cfg = abi.GmcConfig.from_buffer_copy(host.cfg)
    parts = balanced_rows(host, nranks) if rows == "balanced" else round_robin(host.nlat, nranks)
    cfg.lat_ibeg, cfg.lat_iend = parts[rank]
    cfg.rank, cfg.nranks, cfg.device = rank, nranks, device
    return cfg


def band_slice(host, cfg, a, loc):
    """Rows lat_ibeg-hy .. lat_iend+hy (half-lat: .. half_iend+hy) of a global halo'd array, contiguous."""
    hy = host.hy
    half = loc.split("+")[0] in ("lat", "vtx")
    jend = cfg.lat_iend - 1 if (half and cfg.lat_iend == host.nlat) else cfg.lat_iend
    sl = slice(cfg.lat_ibeg - 1, jend + 2 * hy)
    return np.ascontiguousarray(a[sl] if loc == "2d" else a[:, sl])


class GmcoreBand(GmcoreGPU):
    """One latitude band = one handle (what one MPI rank of GMCORE owns)."""

    def __init__(self, host, rank, nranks, device=0, rows="round_robin"):
        self.lib = load_library()
        self.host = host
        self.cfg = band_config(host, rank, nranks, device, rows)
        self.h = C.c_void_p()
        rc = self.lib.gmc_create(C.byref(self.cfg), C.byref(self.h))
        if rc != 0:
            raise RuntimeError("gmc_create failed: " + self.lib.gmc_create_error().decode())
        self._chk(self.lib.gmc_set_mesh(self.h, C.byref(host.mesh_vectors)))
        for which, (ngl, wl, ldl, nga, wa, lda) in enumerate(host.filters):
            self._chk(self.lib.gmc_set_filters(self.h, which, abi.iptr(ngl), abi.dptr(wl), ldl, abi.iptr(nga), abi.dptr(wa), lda))
        self.old, self.new, self.time_step = C.c_int(1), C.c_int(2), C.c_int(0)

    def _b(self, a, loc):
        return None if a is None else band_slice(self.host, self.cfg, a, loc)

    def set_state(self, ic, itime=1):
        self._chk(self.lib.gmc_set_static(self.h, abi.dptr(self._b(ic["gzs"], "2d"))))
        self._chk(self.lib.gmc_upload_state(self.h, itime, abi.dptr(self._b(ic["u_lon"], "lon")), abi.dptr(self._b(ic["v_lat"], "lat")),
                                            abi.dptr(self._b(ic["pt_or_gz"], "cell")), abi.dptr(self._b(ic.get("phs"), "2d"))))

    def set_tracers(self, q, batch=0):
        self.set_tracers_sliced(self.slice_tracers(q), batch)

    # band slicing of global arrays is host set-up; the *_sliced entry points take this rank's rows directly
    # (what a banded Fortran host owns in the first place)
    def slice_state(self, ic):
        loc = {"gzs": "2d", "u_lon": "lon", "v_lat": "lat", "pt_or_gz": "cell", "phs": "2d"}
        return {k: self._b(ic.get(k), loc[k]) for k in loc}

    def slice_tracers(self, q):
        return np.ascontiguousarray(np.stack([self._b(x, "cell") for x in q]))

    def set_state_sliced(self, b, itime=1):
        self._chk(self.lib.gmc_set_static(self.h, abi.dptr(b["gzs"])))
        self._chk(self.lib.gmc_upload_state(self.h, itime, abi.dptr(b["u_lon"]), abi.dptr(b["v_lat"]), abi.dptr(b["pt_or_gz"]), abi.dptr(b["phs"])))

    def set_tracers_sliced(self, qb, batch=0):
        assert qb.shape[0] == self._batch_ntr()[batch]
        self._chk(self.lib.gmc_set_batch_tracers(self.h, batch, abi.dptr(qb)))

    def band_shape(self, loc):
        s = list(self.host.shape(loc))
        half = loc.split("+")[0] in ("lat", "vtx")
        jend = self.cfg.lat_iend - 1 if (half and self.cfg.lat_iend == self.host.nlat) else self.cfg.lat_iend
        s[-2] = jend - self.cfg.lat_ibeg + 1 + 2 * self.host.hy
        return tuple(s)

    def field(self, name, itime=None):
        """Band-local halo'd array of a named field."""
        itime = self.old.value if itime is None else itime
        a = np.zeros(self.band_shape(FIELD_LOC[name]))
        self._chk(self.lib.gmc_get_field(self.h, itime, name.encode(), abi.dptr(a)))
        return a

    def get_tracers(self, batch=0):
        q = np.zeros((self._batch_ntr()[batch],) + self.band_shape("cell"))
        self._chk(self.lib.gmc_get_batch_tracers(self.h, batch, abi.dptr(q)))
        return q

    def download_state(self, itime=None, out=None):
        itime = self.old.value if itime is None else itime
        if out is not None:
            u, v, s, phs = out["u_lon"], out["v_lat"], out["pt_or_gz"], out["phs"]
            for a, loc in ((u, "lon"), (v, "lat"), (s, "cell"), (phs, "2d")):
                if a.shape != tuple(self.band_shape(loc)) or a.dtype != np.float64 or not a.flags.c_contiguous:
                    raise ValueError(f"download_state: out array for '{loc}' must be C-contiguous float64 of shape {tuple(self.band_shape(loc))}")
        else:
            u = np.zeros(self.band_shape("lon")); v = np.zeros(self.band_shape("lat")); s = np.zeros(self.band_shape("cell"))
            phs = np.zeros(self.band_shape("2d"))
        self._chk(self.lib.gmc_download_state(self.h, itime, abi.dptr(u), abi.dptr(v), abi.dptr(s), abi.dptr(phs)))
        return {"u_lon": u, "v_lat": v, "pt_or_gz": s, "phs": phs}

    # one process per GPU: NCCL bootstrap (the host broadcasts the id; MPI_Bcast in GMCORE)
    def nccl_unique_id(self):
        buf = C.create_string_buffer(128)
        if self.lib.gmc_nccl_unique_id(buf) != 0:
            raise RuntimeError("gmc_nccl_unique_id failed (libnccl not loaded)")
        return buf.raw

    def nccl_init(self, uid):
        self._chk(self.lib.gmc_nccl_init(self.h, uid))


def make_band_nccl(host, rank, world, device, rows=None):
    """torch.distributed-launched rank -> its latitude band with the NCCL transport set up.
    rows: "round_robin" (default: the reference's equal row counts; GMCORE_BAND_ROWS overrides) or "balanced"."""
    import torch
    import torch.distributed as dist
    rows = rows or os.environ.get("GMCORE_BAND_ROWS", "round_robin")
    band = GmcoreBand(host, rank, world, device, rows)
    t = torch.zeros(128, dtype=torch.uint8, device=f"cuda:{device}")
    if rank == 0:
        t = torch.frombuffer(bytearray(band.nccl_unique_id()), dtype=torch.uint8).to(f"cuda:{device}")
    dist.broadcast(t, src=0)
    band.nccl_init(bytes(t.cpu().numpy().tobytes()))
    return band


def scatter_rows(host, cfg, band_arr, glob, loc):
    """Copy the owned rows of a band-local halo'd array into the global halo'd array."""
    hy = host.hy
    half = loc.split("+")[0] in ("lat", "vtx")
    jend = cfg.lat_iend - 1 if (half and cfg.lat_iend == host.nlat) else cfg.lat_iend
    n = jend - cfg.lat_ibeg + 1
    if loc == "2d":
        glob[cfg.lat_ibeg - 1 + hy: cfg.lat_ibeg - 1 + hy + n] = band_arr[hy:hy + n]
    else:
        glob[:, cfg.lat_ibeg - 1 + hy: cfg.lat_ibeg - 1 + hy + n] = band_arr[:, hy:hy + n]


class GmcoreBandGroup:
    """All latitude bands in ONE process on ONE device (gmc_group_link): the single-GPU stand-in for
    the N-rank run, used to check that results do not depend on the band count."""

    def __init__(self, host, nbands, device=0, rows="round_robin"):
        self.host = host
        self.bands = [GmcoreBand(host, r, nbands, device, rows) for r in range(nbands)]
        self.lib = self.bands[0].lib
        arr = (C.c_void_p * nbands)(*[b.h for b in self.bands])
        if self.lib.gmc_group_link(arr, nbands) != 0:
            self.bands[0]._chk(1)
        self.lead = self.bands[0]

    def set_state(self, ic):
        for b in self.bands:
            b.set_state(ic)

    def add_tracer_batch(self, ntracers, dt_adv):
        return [b.add_tracer_batch(ntracers, dt_adv) for b in self.bands][0]

    def set_tracers(self, q, batch=0):
        for b in self.bands:
            b.set_tracers(q, batch)

    def prepare_run(self):
        self.lead.prepare_run()

    def run_steps(self, n, diag=True):
        d = self.lead.run_steps(n, diag)
        for b in self.bands[1:]:
            b.old.value, b.new.value, b.time_step.value = self.lead.old.value, self.lead.new.value, self.lead.time_step.value
        return d

    def diagnose(self):
        return self.lead.diagnose()

    def field(self, name):
        loc = FIELD_LOC[name]
        g = np.zeros(self.host.shape(loc))
        for b in self.bands:
            scatter_rows(self.host, b.cfg, b.field(name), g, loc)
        return g

    def get_tracers(self, batch=0):
        g = np.zeros((self.lead._batch_ntr()[batch],) + self.host.shape("cell"))
        for b in self.bands:
            qb = b.get_tracers(batch)
            for l in range(g.shape[0]):
                scatter_rows(self.host, b.cfg, qb[l], g[l], "cell")
        return g

    def close(self):
        for b in self.bands:
            b.close()


def band_parity(rank, world, device, nlon=192, nlat=97, nlev=16, dt=300.0, nsteps=6, case="baroclinic_wave"):
    """The reference's serial-vs-parallel check for the NCCL path: this rank's latitude band of a `world`-rank run
    (stage exchange over ncclSend/ncclRecv, overlapped with the interior filter rows unless GMCORE_NO_OVERLAP=1)
    against ONE band owning the sphere on the same device, on this rank's own rows.  Needs an initialised
    torch.distributed NCCL group.  Returns {"worst_rel_l2", "diag"} (max over ranks); the bar is 1e-13."""
    import torch
    import torch.distributed as dist
    from .host import Host
    host = Host(nlon, nlat, nlev, dt, case)
    ic = host.initial_condition()
    q = np.zeros((host.cfg.ntracers,) + host.shape("cell"))
    kk, jj, ii = np.meshgrid(np.arange(q.shape[1]), np.arange(q.shape[2]), np.arange(q.shape[3]), indexing="ij")
    q[0] = 1.0 + 0.5 * np.sin(2 * np.pi * ii / q.shape[3]) * np.cos(np.pi * (jj / q.shape[2] - 0.5)) + 0.01 * kk
    band = make_band_nccl(host, rank, world, device)
    band.set_state(ic); band.set_tracers(q); band.prepare_run()
    db = band.run_steps(nsteps)
    one = GmcoreGPU(host, device=device)
    one.set_state(ic); one.set_tracers(q); one.prepare_run()
    d1 = one.run_steps(nsteps)
    hy, hx = host.hy, host.hx
    worst = 0.0
    for name, loc in (("u_lon", "lon"), ("v_lat", "lat"), ("pt", "cell"), ("phs", "2d"), ("u_f", "lon"), ("v_f", "lat")):
        a = band.field(name); b = band_slice(host, band.cfg, one.field(name), loc)
        if loc == "2d":
            a = a.reshape(a.shape[-2:]); b = b.reshape(b.shape[-2:])
        a = a[..., hy:a.shape[-2] - hy, hx:-hx]; b = b[..., hy:b.shape[-2] - hy, hx:-hx]
        if loc != "2d":
            a = a[2:2 + nlev]; b = b[2:2 + nlev]
        worst = max(worst, float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300)))
    qa = band.get_tracers()[0]; qb = band_slice(host, band.cfg, one.get_tracers()[0], "cell")
    sl = (slice(2, 2 + nlev), slice(hy, -hy), slice(hx, -hx))
    worst = max(worst, float(np.linalg.norm((qa - qb)[sl].ravel()) / np.linalg.norm(qb[sl].ravel())))
    dd = float(np.abs(db[:, [0, 1, 7]] / d1[:, [0, 1, 7]] - 1).max())
    band.close(); one.close()
    t = torch.tensor([worst, dd], dtype=torch.float64, device=f"cuda:{device}")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {"worst_rel_l2": float(t[0].item()), "diag": float(t[1].item()),
            "case": f"{case} {nlon}x{nlat}x{nlev}, {nsteps} steps, {world} NCCL bands vs 1 band, overlap "
                    f"{'off' if os.environ.get('GMCORE_NO_OVERLAP') == '1' else 'on'}"}
