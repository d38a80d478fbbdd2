"""The CUDA path (through the C ABI) against analytic steady states -- the cases and bounds of
tests/test_known_answers_cpu.py, where the oracle is held to them.  Here no oracle is involved at all: Williamson
case 2 (tests/swm/steady_geostrophic_flow_test_mod.F90 of the reference) must converge at second order and the
Jablonowski-Williamson steady state (tests/steady_state_test_mod.F90) must stay zonally symmetric to round-off and
balanced to truncation error.  One finer mesh than on the CPU is added for case 2 (240 x 120: a third point on the
convergence line)."""
import pytest

from test_known_answers_cpu import JW_COARSE, JW_FINE, check_jw_steady, check_w2, run_jw_steady, run_w2

pytestmark = pytest.mark.gpu


def test_gpu_holds_williamson_case_2(gpu_required):
    from gmcore_b200.model import GmcoreGPU
    r60, r120, r240 = (run_w2(GmcoreGPU, n, n // 2, dt) for n, dt in ((60, 600.0), (120, 300.0), (240, 150.0)))
    check_w2(r60, r120)
    assert 3.0 < r120["l2"] / r240["l2"] < 5.5 and r240["l2"] < 1.5e-5, (r120["l2"], r240["l2"])
    assert r240["asym_gz"] < 1e-12 and r240["asym_u"] < 1e-11 and r240["tm_drift"] < 1e-13, r240


def test_gpu_holds_the_jw_steady_state(gpu_required):
    from gmcore_b200.model import GmcoreGPU
    check_jw_steady(run_jw_steady(GmcoreGPU, *JW_COARSE), run_jw_steady(GmcoreGPU, *JW_FINE))
