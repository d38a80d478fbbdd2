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


@pytest.mark.parametrize("name", ["deform_case4", "moving_vortices", "dcmip12"])
def test_adv_driver_cases_gpu_match_the_oracle(gpu_required, name):
    """The reference's other stand-alone advection cases through gmc_adv_prescribe(_we) / gmc_adv_step:
    time-dependent prescribed winds, four / two tracers, and for DCMIP 1-2 a prescribed vertical mass flux.  Per-step
    relative L2 <= 1e-12 against the oracle on every tracer (the oracle is held to the analytic properties of each
    case in tests/test_advection_cpu.py)."""
    import numpy as np
    from adv_cases import DeformCase4, Dcmip12, MovingVortices
    from gmcore_b200.model import GmcoreGPU
    from oracle.orc import Oracle
    case = {"deform_case4": lambda: DeformCase4(90, 46, 1, 3600.0), "moving_vortices": lambda: MovingVortices(90, 46, 1, 3600.0),
            "dcmip12": lambda: Dcmip12(90, 46, 16, 1800.0)}[name]()
    host = case.host
    q = case.tracers()
    gpu = GmcoreGPU(host); orc = Oracle(host)
    gpu.set_tracers(q); orc.set_tracers(q)
    for s in range(12):
        u, v, we = case.winds((s + 1) * case.dt)
        for m in (gpu, orc):
            m.adv_prescribe(u, v, we_lev=we)
            m.adv_step()
        qg, qo = gpu.get_tracers(), orc.get_tracers()
        for l in range(q.shape[0]):
            a = host.interior(qg[l], "cell"); b = host.interior(qo[l], "cell")
            e = np.linalg.norm(a - b) / np.linalg.norm(b)
            assert e <= 1e-12, (name, s, l, e)
    gpu.close()
