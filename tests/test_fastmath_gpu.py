"""The device build of pow_pos / log_pos returns the host build's bits (gmc_fastmath_eval through the C ABI against
oracle/_build/libfastmath_host.so), so the < 1 ulp bounds of tests/test_fastmath_cpu.py hold on the GPU; and the fields
the routines feed (t, gz_lev, ph_exn_lev: calc_t / calc_gz_lev / calc_ph, operators_mod.F90:136-190,433-469) stay
within the parity bar of the oracle, which uses glibc's pow / log."""
import ctypes as C

import numpy as np
import pytest

from fastmath_util import KAPPA, dptr, host_lib, samples

pytestmark = pytest.mark.gpu


def test_device_results_are_the_host_results(gpu_required):
    from gmcore_b200.model import load_library
    dev = load_library(); host = host_lib()
    x = samples(100000)
    with np.errstate(all="ignore"):
        x = np.ascontiguousarray(np.concatenate([x, [0.0, -1.0, np.inf, 5e-324, 1e-310]]))
    a = np.empty_like(x); b = np.empty_like(x)
    for what, y in ((0, KAPPA), (0, 2.0 / 7.0), (0, -0.7), (0, 2.5), (1, 0.0)):
        assert dev.gmc_fastmath_eval(what, dptr(x), y, dptr(a), x.size) == 0
        host.fm_eval(what, dptr(x), y, dptr(b), x.size)
        inside = np.isfinite(x) & (x >= 2.2250738585072014e-308)
        if what == 0:
            with np.errstate(all="ignore"):
                inside &= np.abs(y * np.log(np.where(inside, x, 1.0))) < 690.0
        # inside the domain: bit for bit.  Outside it both sides defer to their own library (CUDA's / glibc's).
        np.testing.assert_array_equal(a[inside].view(np.int64), b[inside].view(np.int64), err_msg=f"what={what} y={y}")
        out = ~inside & np.isfinite(b)
        if out.any():
            np.testing.assert_allclose(a[out], b[out], rtol=1e-15, atol=1e-300)      # (subnormal results: the two libraries differ)
