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
               "gmc_synchronize", "gmc_launch_count", "gmc_last_error"]


def load_library():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = os.path.join(_LIBDIR, "libgmcore_b200.so")
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
    lib.gmc_prepare_run.argtypes = [V, C.c_int]
    lib.gmc_time_integrate.argtypes = [V, C.c_int, C.c_int, C.c_double, C.c_int]
    lib.gmc_run_steps.argtypes = [V, I, I, I, C.c_int, P]
    lib.gmc_diagnose.argtypes = [V, C.c_int, P]
    lib.gmc_get_field.argtypes = [V, C.c_int, C.c_char_p, P]
    lib.gmc_synchronize.argtypes = [V]
    lib.gmc_launch_count.argtypes = [V, C.POINTER(C.c_longlong)]
    lib.gmc_last_error.argtypes = [V, C.c_char_p, C.c_int]
    lib.gmc_struct_sizes.argtypes = [I, I]
    lib.gmc_timer_record.argtypes = [V, C.c_int]
    lib.gmc_timer_elapsed_ms.argtypes = [V, P]
    lib.gmc_profile.argtypes = [V, C.c_int]
    lib.gmc_profile_report.argtypes = [V, C.c_char_p, C.c_int]
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
    "wedudlev": "lon", "wedvdlev": "lat", "dphs": "2d",
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

    def set_tracers(self, q):
        self._chk(self.lib.gmc_set_tracers(self.h, abi.dptr(q)))

    def get_tracers(self):
        q = np.zeros((self.host.cfg.ntracers,) + self.host.shape("cell"))
        self._chk(self.lib.gmc_get_tracers(self.h, abi.dptr(q)))
        return q

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
    def download_state(self, itime=None):
        itime = self.old.value if itime is None else itime
        hst = self.host
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
