"""gmc_diagnose_static: `diagnose` with its scratch in the library's static device array (no allocation).

The entry point has no counterpart in the reference (include/gmcore_gpu.h says why it exists); its contract is that it
returns what gmc_diagnose returns.  Both launch the same two kernels (k_diag_partial, k_diag_final) in the same
order, so the sums agree to the last bit; they are held to the 1e-13 the other diagnostics tests use.

Checked through the C ABI: two calls in a row on one handle (the state moves in between), then a second handle of a
different size and core (the array still holds the first handle's block sums), then the two refusals (a band handle,
a grid beyond GMC_STATIC_SCRATCH_DOUBLES).  This file sorts last on purpose.
"""
import numpy as np
import pytest

from util import assert_diag_close, smooth_tracer

pytestmark = pytest.mark.gpu
RTOL = 1e-13


def gpu_model(nlon, nlat, nlev, dt, case, **kw):
    from gmcore_b200.host import Host
    from gmcore_b200.model import GmcoreGPU
    host = Host(nlon, nlat, nlev, dt, case, **kw)
    gpu = GmcoreGPU(host)
    gpu.set_state(host.initial_condition())
    if host.cfg.ntracers:
        gpu.set_tracers(smooth_tracer(host, host.cfg.ntracers))
    gpu.prepare_run()
    return host, gpu


def same_sums(gpu, what):
    want = gpu.diagnose()
    got = gpu.diagnose_static()
    assert np.all(np.isfinite(got)) and got[0] > 0, (what, got)
    assert_diag_close(got, want, rtol=RTOL)
    again = gpu.diagnose_static()          # the array now holds this very call's leftovers
    np.testing.assert_array_equal(again, got, err_msg=what)
    return got


def test_static_diagnose_matches_diagnose_twice_in_a_row(gpu_required):
    host, gpu = gpu_model(96, 49, 8, 450.0, "baroclinic_wave")
    d0 = same_sums(gpu, "hydro, initial state")
    gpu.run_steps(3)
    d1 = same_sums(gpu, "hydro, after 3 steps")
    assert d1[1] != d0[1]                  # the second call saw the new state, not the first call's sums
    # a smaller shallow-water handle next: fewer block sums than the array still holds from the calls above
    host2, swm = gpu_model(72, 37, 1, 300.0, "rossby_haurwitz_wave")
    same_sums(swm, "swm, initial state")
    swm.run_steps(2)
    same_sums(swm, "swm, after 2 steps")
    # and back to the first handle: its state has not moved, so neither have its sums
    np.testing.assert_array_equal(gpu.diagnose_static(), d1)
    swm.close(); gpu.close()


def test_static_diagnose_refuses_bands_and_oversized_grids(gpu_required):
    from gmcore_b200.host import Host
    from gmcore_b200.model import GmcoreBandGroup, GmcoreGPU
    host = Host(64, 33, 8, 600.0, "baroclinic_wave")
    grp = GmcoreBandGroup(host, 2)
    grp.set_state(host.initial_condition())
    with pytest.raises(RuntimeError, match="single-band"):
        grp.lead.diagnose_static(1)
    grp.close()
    # 32 + 7 * 1201 rows * 64 levels = 538080 doubles > GMC_STATIC_SCRATCH_DOUBLES (524288): refused before any launch
    tall = Host(32, 1201, 64, 1.0, "baroclinic_wave", ntracers=0)
    big = GmcoreGPU(tall)
    with pytest.raises(RuntimeError, match="GMC_STATIC_SCRATCH_DOUBLES"):
        big.diagnose_static(1)
    big.close()
