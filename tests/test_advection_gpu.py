"""The stand-alone advection test of the reference (solid rotation over the poles, see test_advection_cpu.py)
on the CUDA path: parity with the oracle after every 48th step and the same analytic error norms."""
import numpy as np
import pytest

from test_advection_cpu import error_norms, solid_rotation_setup

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("flux", [0, 1])
def test_solid_rotation_gpu_matches_oracle_and_analytic(gpu_required, flux):
    """flux 0 = PPM, 1 = van Leer (CFL > 1 in the polar rows exercises the whole-cell sums of both)."""
    from gmcore_b200.model import GmcoreGPU
    from oracle.orc import Oracle
    host, u, v, q, bell, area, nsteps = solid_rotation_setup(ffsl_flux_type=flux)
    gpu = GmcoreGPU(host); orc = Oracle(host)
    gpu.set_tracers(q); orc.set_tracers(q)
    for n in range(nsteps + 1):
        gpu.adv_prescribe(u, v); gpu.adv_step()
        orc.adv_prescribe(u, v); orc.adv_step()
        if n % 48 == 0 or n == nsteps:
            a = host.interior(gpu.get_tracers()[1], "cell"); b = host.interior(orc.get_tracers()[1], "cell")
            e = np.linalg.norm((a - b).ravel()) / np.linalg.norm(b.ravel())
            assert e <= 1e-12, (n, e)
    qn = gpu.get_tracers()
    l1, l2, linf, mass, qmin = error_norms(host, qn, bell, area)
    assert (l2 < 0.25 and l1 < 0.35 and linf < 0.35) if flux == 0 else l2 < 0.6, (l1, l2, linf)
    assert abs(mass) < 1e-12 and qmin >= 0.0, (mass, qmin)
    q0 = host.interior(qn[0], "cell")
    assert np.abs(q0 - 1).max() < 1e-3
    gpu.close()
