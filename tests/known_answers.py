"""Initial states with a known exact answer, built in Python on the stand-in host's mesh.

These are checks that do not go through a second restatement of the reference: the answer is an analytic steady state
of the continuous equations, so a model (oracle or CUDA) that keeps it to O(dx^2) and keeps its symmetry to round-off is
right for reasons the oracle and the kernels do not share.

  steady_geostrophic_flow   Williamson et al. (1992) case 2, alpha = 0: the reference's
                            tests/swm/steady_geostrophic_flow_test_mod.F90:20-66 (u0 = 2 pi a / 12 days, gz0 = 2.94e4)
  steady_state              the Jablonowski-Williamson balanced state without the wind perturbation, the reference's
                            tests/steady_state_test_mod.F90 (same analytic fields as baroclinic_wave_test_mod.F90 minus u')
"""
import numpy as np

GZ0 = 2.94e4


def steady_geostrophic_flow(host, ic, alpha=0.0):
    """Overwrites the interior of a shallow-water initial state with Williamson case 2 (gzs = 0)."""
    a, om = host.cfg.radius, host.cfg.omega
    u0 = 2.0 * np.pi * a / (12.0 * 86400.0)
    hx, hy = host.hx, host.hy
    fl = host.vec["full_lon"][hx:hx + host.nlon]; hl = host.vec["half_lon"][hx:hx + host.nlon]
    cf = host.vec["full_cos_lat"][hy:hy + host.nlat]; sf = host.vec["full_sin_lat"][hy:hy + host.nlat]
    ca, sa = np.cos(alpha), np.sin(alpha)
    u = host.interior(ic["u_lon"], "lon"); v = host.interior(ic["v_lat"], "lat"); gz = host.interior(ic["pt_or_gz"], "cell")
    u[...] = (u0 * (cf[:, None] * ca + np.cos(hl)[None, :] * sf[:, None] * sa))[None]
    # (the reference's v uses full_cos_lon where sin(lon) is meant, :50; it multiplies sin(alpha) = 0 there)
    v[...] = (-u0 * np.sin(fl) * sa)[None, None, :]
    gz[...] = (GZ0 - (a * om * u0 + 0.5 * u0 ** 2) * (sf[:, None] * ca - np.cos(fl)[None, :] * cf[:, None] * sa) ** 2)[None]
    ic["gzs"][...] = 0.0
    return ic


def zonal_asymmetry(host, a, loc):
    """max |a - zonal mean| / max |a| over the interior."""
    x = host.interior(a, loc)
    return float(np.abs(x - x.mean(axis=-1, keepdims=True)).max() / max(np.abs(x).max(), 1e-300))


def l2_error(host, a, a0, loc="cell"):
    """Area-weighted normalised L2 error of Williamson et al. (1992) eq. 83 on the cell grid."""
    w = host.vec["area_cell"][host.hy:host.hy + host.nlat][None, :, None]
    x = host.interior(a, loc); x0 = host.interior(a0, loc)
    return float(np.sqrt((w * (x - x0) ** 2).sum() / (w * x0 ** 2).sum()))


def jw_steady_state(host, ic):
    """Removes the zonal wind perturbation of the baroclinic-wave state: every longitude gets the column of the
    longitude opposite the perturbation centre (20E), where exp(-(r / (a/10))^2) underflows to 0."""
    u = host.interior(ic["u_lon"], "lon")
    lon = host.vec["half_lon"][host.hx:host.hx + host.nlon]
    far = int(np.argmin(np.abs(lon - (np.pi / 9 + np.pi))))
    u[...] = u[:, :, far:far + 1]
    return ic
