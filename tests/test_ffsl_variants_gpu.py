"""The FFSL kernel variants are bit-identical to one another.

run_ffsl_h (gmcore_b200/csrc/hydro.cuh) picks, by environment variable read when the handle is created:
  default                 warp-private x sweep + register-marching y sweep (inner), fused marching kernel (outer)
  GMCORE_OLD_HFLX=1       the round-1 per-edge kernels (k_hflx + k_ffsl_q)
  GMCORE_FFSL_FUSED=1     fused marching kernel for the inner call too
  GMCORE_FFSL_FUSED=0     split x / y kernels for both calls
  ... + GMCORE_FFSL_XBLOCK=1   block-staged x sweep instead of the warp-private one
All of them restate ffsl_mod.F90:354-406 / :778-881 with the same expressions in the same order, so every variant
must reproduce the default's fields to the last bit.  Each variant runs in its own process (the switch is read once)
and prints a SHA-256 of the prognostic fields + tracers after a few steps; the default's parity with the oracle is
what every other GPU test checks.
"""
import hashlib
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

VARIANTS = {
    "default": {},
    "round1": {"GMCORE_OLD_HFLX": "1"},
    "fused": {"GMCORE_FFSL_FUSED": "1"},
    "split": {"GMCORE_FFSL_FUSED": "0"},
    "split_xblock": {"GMCORE_FFSL_FUSED": "0", "GMCORE_FFSL_XBLOCK": "1"},
}

CHILD = r"""
import hashlib, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, %(root)r + "/tests")
import numpy as np
from gmcore_b200.host import Host
from gmcore_b200.model import GmcoreGPU
from util import smooth_tracer
from test_baseline_sizes_gpu import cross_polar_flow
h = hashlib.sha256()
for (nlon, nlat, nlev, dt, polar) in ((96, 49, 8, 450.0, False), (192, 97, 6, 300.0, True)):
    host = Host(nlon, nlat, nlev, dt, "baroclinic_wave")
    ic = host.initial_condition()
    if polar:
        cross_polar_flow(host, ic)          # |cflx| >= 1 on the polar rows: the gather path of every variant
    gpu = GmcoreGPU(host); gpu.set_state(ic)
    gpu.set_tracers(smooth_tracer(host, host.cfg.ntracers))
    gpu.prepare_run()
    gpu.run_steps(3)
    for name in ("u_lon", "v_lat", "pt", "phs"):
        h.update(np.ascontiguousarray(gpu.field(name)).tobytes())
    h.update(np.ascontiguousarray(gpu.get_tracers(0)).tobytes())
    gpu.close()
print("DIGEST", h.hexdigest())
"""


def digest(env_extra):
    env = dict(os.environ)
    for k in ("GMCORE_OLD_HFLX", "GMCORE_FFSL_FUSED", "GMCORE_FFSL_XBLOCK"):
        env.pop(k, None)
    env.update(env_extra)
    out = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("DIGEST ")]
    assert lines, out.stdout[-2000:]
    return lines[-1].split()[1]


def test_ffsl_variants_are_bit_identical(gpu_required):
    ref = digest(VARIANTS["default"])
    bad = {}
    for name, env in VARIANTS.items():
        if name == "default":
            continue
        d = digest(env)
        if d != ref:
            bad[name] = d
    assert not bad, f"variants differ from the default ({ref}): {bad}"
