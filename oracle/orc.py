"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes binding of oracle/_build/liboracle*.so.

Importers allowed: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline /
--impl reference legs.  The product package (gmcore_b200/) never imports this.
Parity unpinned by the reference's own tests (it ships none) — see orc_core.hpp.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)


def build():
    subprocess.check_call(["make", "-C", _ROOT, "oracle"], stdout=subprocess.DEVNULL)


def _load(fast):
    path = os.path.join(_HERE, "_build", "liboracle_fast.so" if fast else "liboracle.so")
    if not os.path.exists(path):
        build()
    lib = C.CDLL(path)
    P = C.POINTER(C.c_double)
    lib.orc_create.restype = C.c_void_p
    lib.orc_create.argtypes = [C.c_void_p, C.c_void_p, P]
    lib.orc_destroy.argtypes = [C.c_void_p]
    lib.orc_set_filters.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), P, C.c_int, C.POINTER(C.c_int), P, C.c_int]
    lib.orc_set_static.argtypes = [C.c_void_p, P]
    lib.orc_upload_state.argtypes = [C.c_void_p, C.c_int, P, P, P, P]
    lib.orc_set_tracers.argtypes = [C.c_void_p, P]
    lib.orc_get_tracers.argtypes = [C.c_void_p, P]
    lib.orc_add_tracer_batch.argtypes = [C.c_void_p, C.c_int, C.c_double]
    lib.orc_set_batch_tracers.argtypes = [C.c_void_p, C.c_int, P]
    lib.orc_get_batch_tracers.argtypes = [C.c_void_p, C.c_int, P]
    lib.orc_adv_prescribe.argtypes = [C.c_void_p, C.c_int, P, P]
    lib.orc_adv_prescribe_we.argtypes = [C.c_void_p, C.c_int, P]
    lib.orc_adv_step.argtypes = [C.c_void_p]
    lib.orc_interp_pressure_level.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_double, C.c_int, C.c_double, P]
    lib.orc_prepare_run.argtypes = [C.c_void_p]
    lib.orc_run_steps.argtypes = [C.c_void_p, C.c_int, P]
    lib.orc_diagnose.argtypes = [C.c_void_p, C.c_int, P]
    for f in ("orc_old_index", "orc_new_index", "orc_time_step"):
        getattr(lib, f).argtypes = [C.c_void_p]
    lib.orc_get_field.restype = C.c_longlong
    lib.orc_get_field.argtypes = [C.c_void_p, C.c_int, C.c_char_p, P]
    lib.orc_field_shape.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.POINTER(C.c_int)]
    lib.orc_last_error.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
    return lib


def _dp(a):
    if a is None:
        return C.POINTER(C.c_double)()
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_double))


class Oracle:
    """CPU restatement driven with the same inputs as the GPU library.

    `host` is a gmcore_b200.host.Host (mesh vectors, filters, config); only plain
    numpy arrays / ctypes structs are taken from it.
    """

    def __init__(self, host, fast=False):
        self.lib = _load(fast)
        self.host = host
        self.h = self.lib.orc_create(C.addressof(host.cfg), C.addressof(host.mesh_vectors), _dp(host.vec["full_lat_deg"]))
        self.batch_ntr = {0: host.cfg.ntracers} if host.cfg.ntracers > 0 else {}
        for which, (ngl, wl, ldl, nga, wa, lda) in enumerate(host.filters):
            self.lib.orc_set_filters(self.h, which, ngl.ctypes.data_as(C.POINTER(C.c_int)), _dp(wl), ldl,
                                     nga.ctypes.data_as(C.POINTER(C.c_int)), _dp(wa), lda)

    def set_state(self, ic, itime=1):
        self.lib.orc_set_static(self.h, _dp(ic["gzs"]))
        self.lib.orc_upload_state(self.h, itime, _dp(ic["u_lon"]), _dp(ic["v_lat"]), _dp(ic["pt_or_gz"]), _dp(ic.get("phs")))

    def add_tracer_batch(self, ntracers, dt_adv):
        """adv_add_tracer + adv_allocate_tracers (adv_mod.F90:314-418); returns the batch index."""
        b = self.lib.orc_add_tracer_batch(self.h, ntracers, float(dt_adv))
        self.batch_ntr[b] = ntracers
        return b

    def set_tracers(self, q, batch=0):
        assert q.shape[0] == self.batch_ntr[batch]
        self.lib.orc_set_batch_tracers(self.h, batch, _dp(q))

    def get_tracers(self, batch=0):
        q = np.zeros((self.batch_ntr[batch],) + self.host.shape("cell"))
        self.lib.orc_get_batch_tracers(self.h, batch, _dp(q))
        return q

    # advection-only driver (gmcore_adv_driver.F90:84-88)
    def adv_prescribe(self, u_f, v_f, itime=None, we_lev=None):
        it = self.new_index if itime is None else itime
        self.lib.orc_adv_prescribe(self.h, it, _dp(u_f), _dp(v_f))
        if we_lev is not None:
            self.lib.orc_adv_prescribe_we(self.h, it, _dp(we_lev))

    def adv_step(self):
        self._check(self.lib.orc_adv_step(self.h))

    def interp_pressure_level(self, var, po, logp=False, fill=np.nan, itime=None):
        y = np.zeros(self.host.shape("2d"))
        self.lib.orc_interp_pressure_level(self.h, self.old_index if itime is None else itime, var.encode(), float(po), int(logp), float(fill), _dp(y))
        return y

    def prepare_run(self):
        self._check(self.lib.orc_prepare_run(self.h))

    def run_steps(self, n):
        diag = np.zeros((n, 8))
        self._check(self.lib.orc_run_steps(self.h, n, _dp(diag)))
        return diag

    def diagnose(self, itime=None):
        out = np.zeros(8)
        self.lib.orc_diagnose(self.h, self.old_index if itime is None else itime, _dp(out))
        return out

    @property
    def old_index(self):
        return self.lib.orc_old_index(self.h)

    @property
    def new_index(self):
        return self.lib.orc_new_index(self.h)

    def field(self, name, itime=None):
        itime = self.old_index if itime is None else itime
        shp = (C.c_int * 6)()
        if self.lib.orc_field_shape(self.h, itime, name.encode(), shp) != 0:
            raise KeyError(name)
        a = np.zeros((shp[2], shp[1], shp[0]))
        self.lib.orc_get_field(self.h, itime, name.encode(), _dp(a))
        return a

    def _check(self, rc):
        if rc != 0:
            buf = C.create_string_buffer(512)
            self.lib.orc_last_error(self.h, buf, 512)
            raise RuntimeError("oracle: " + buf.value.decode())

    def __del__(self):
        try:
            self.lib.orc_destroy(self.h)
        except Exception:
            pass
