"""Shared helpers for the parity tests."""
import numpy as np

from gmcore_b200.host import Host


def rel_l2(a, b):
    """Relative L2 difference |a-b| / |b| over the given (interior) arrays."""
    d = np.linalg.norm((a - b).ravel())
    n = np.linalg.norm(b.ravel())
    return d / n if n > 0 else d


PROGNOSTIC_SWM = [("u_lon", "lon"), ("v_lat", "lat"), ("gz", "cell")]
PROGNOSTIC_HYDRO = [("u_lon", "lon"), ("v_lat", "lat"), ("pt", "cell"), ("phs", "2d")]


def make_pair(nlon, nlat, nlev, dt, case, **kw):
    """Host + oracle + GPU model on identical inputs."""
    from gmcore_b200.model import GmcoreGPU
    from oracle.orc import Oracle
    host = Host(nlon, nlat, nlev, dt, case, **kw)
    ic = host.initial_condition()
    orc = Oracle(host)
    orc.set_state(ic)
    gpu = GmcoreGPU(host)
    gpu.set_state(ic)
    return host, ic, orc, gpu


def assert_diag_close(dg, do, rtol=1e-13, tav_atol=1e-6):
    """diag = tm, te, tav, tpe, te_ke, te_ie, te_pe, tpt.  tav is a signed sum that
    cancels to ~0 (|terms| ~ 1e2 each), so it is compared absolutely."""
    dg = np.atleast_2d(dg); do = np.atleast_2d(do)
    # tm, te (the conservation criteria): rtol; the other integrals are compared at
    # 10*rtol: the oracle sums ~1e4..1e7 terms sequentially, the GPU as a tree.
    np.testing.assert_allclose(dg[:, [0, 1]], do[:, [0, 1]], rtol=2 * rtol, atol=1e-30)
    idx = [3, 7]
    np.testing.assert_allclose(dg[:, idx], do[:, idx], rtol=10 * rtol, atol=1e-30)
    # energy components: te_pe of the hydrostatic core is a signed, cancelling sum
    # (gzs changes sign), so the parts are compared relative to the total energy.
    for q in (4, 5, 6):
        assert np.abs(dg[:, q] - do[:, q]).max() <= 10 * rtol * np.abs(do[:, 1]).max(), q
    assert np.abs(dg[:, 2] - do[:, 2]).max() < tav_atol


def smooth_tracer(host, n):
    """Positive, smooth tracer fields in the halo'd host layout (interior written)."""
    q = np.zeros((n,) + host.shape("cell"))
    lon = host.vec["full_lon"][host.hx:host.hx + host.nlon]
    lat = host.vec["full_lat"][host.hy:host.hy + host.nlat]
    for l in range(n):
        f = 1.0 + 0.5 * np.cos(lat)[:, None] ** 2 * np.sin((l + 2) * lon)[None, :] + 0.3 * np.sin(lat)[:, None]
        for k in range(host.nlev):
            host.interior(q[l], "cell")[k] = f * (1.0 + 0.1 * k / host.nlev)
    return q
