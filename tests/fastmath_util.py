"""ctypes access to the host build of fastmath.cuh (oracle/_build/libfastmath_host.so, tests/fastmath_host.cpp) and the
argument sets of the fastmath tests."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KAPPA = 287.04 / 1004.64


def host_lib():
    lib = C.CDLL(os.path.join(ROOT, "oracle/_build/libfastmath_host.so"))
    P = C.POINTER(C.c_double)
    lib.fm_eval.argtypes = [C.c_int, P, C.c_double, P, C.c_int]
    lib.fm_libm.argtypes = [C.c_int, P, C.c_double, P, C.c_int]
    lib.fm_max_ulp.argtypes = [C.c_int, P, C.c_double, P, C.c_int]
    lib.fm_max_ulp.restype = C.c_double
    return lib


def dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def samples(n, seed=7):
    """Arguments the kernels produce (interface pressures, ph / p0, interface pressure ratios) and a wide sweep,
    plus the edges of the mantissa reduction (powers of two, sqrt(2) 2^k, neighbours of 1)."""
    r = np.random.default_rng(seed)
    x = np.concatenate([
        200.0 + r.random(n) * 110000.0,
        2.0e-3 + r.random(n) * 1.1,
        1.0 + (r.random(n) - 0.3) * 0.6,
        np.exp((r.random(n) - 0.5) * 60.0),
        np.exp((r.random(n) - 0.5) * 1400.0),                 # the whole normal range
    ])
    edge = []
    for k in (-1000, -60, -1, 0, 1, 17, 900):
        for m in (1.0, np.sqrt(2.0), np.nextafter(np.sqrt(2.0), 2), np.nextafter(np.sqrt(2.0), 1), np.nextafter(1.0, 2), np.nextafter(1.0, 0), np.nextafter(2.0, 1)):
            edge.append(np.ldexp(m, k))
    return np.ascontiguousarray(np.concatenate([x, np.array(edge)]))
