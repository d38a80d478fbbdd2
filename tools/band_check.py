"""torchrun -n N tools/band_check.py: the NCCL latitude-band run (one process per GPU, stage exchange overlapped
with interior work) against the same case on ONE band, on this rank's own rows.  Bar: rel L2 <= 1e-13."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
from gmcore_b200.host import Host
from gmcore_b200.model import GmcoreGPU, make_band_nccl, band_slice

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nlon, nlat, nlev, dt, nsteps = 192, 97, 16, 300.0, 6
host = Host(nlon, nlat, nlev, dt, "baroclinic_wave")
ic = host.initial_condition()
q = np.zeros((host.cfg.ntracers,) + host.shape("cell"))
kk, jj, ii = np.meshgrid(np.arange(q.shape[1]), np.arange(q.shape[2]), np.arange(q.shape[3]), indexing="ij")
q[0] = 1.0 + 0.5 * np.sin(2 * np.pi * ii / q.shape[3]) * np.cos(np.pi * (jj / q.shape[2] - 0.5)) + 0.01 * kk
band = make_band_nccl(host, rank, world, local)
band.set_state(ic); band.set_tracers(q); band.prepare_run()
db = band.run_steps(nsteps)
one = GmcoreGPU(host, device=local)
one.set_state(ic); one.set_tracers(q); one.prepare_run()
d1 = one.run_steps(nsteps)
hy = host.hy
worst = 0.0
for name, loc in (("u_lon", "lon"), ("v_lat", "lat"), ("pt", "cell"), ("phs", "2d"), ("u_f", "lon"), ("v_f", "lat")):
    a = band.field(name); b = band_slice(host, band.cfg, one.field(name), loc)
    if loc == "2d":
        a = a.reshape(a.shape[-2:]); b = b.reshape(b.shape[-2:])
    a = a[..., hy:a.shape[-2] - hy, host.hx:-host.hx]; b = b[..., hy:b.shape[-2] - hy, host.hx:-host.hx]
    if loc != "2d":
        a = a[2:2 + nlev]; b = b[2:2 + nlev]
    e = np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300)
    worst = max(worst, e)
qa = band.get_tracers()[0]; qb = band_slice(host, band.cfg, one.get_tracers()[0], "cell")
e = np.linalg.norm((qa - qb)[2:2 + nlev, hy:-hy, host.hx:-host.hx].ravel()) / np.linalg.norm(qb[2:2 + nlev, hy:-hy, host.hx:-host.hx].ravel())
worst = max(worst, e)
dd = np.abs(db[:, [0, 1, 7]] / d1[:, [0, 1, 7]] - 1).max()
t = torch.tensor([worst, dd], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    ok = t[0].item() <= 1e-13 and t[1].item() <= 1e-13
    print(f"band_check N={world} overlap={'off' if os.environ.get('GMCORE_NO_OVERLAP') else 'on'}: worst rel L2 {t[0].item():.2e}, diag {t[1].item():.2e} -> {'OK' if ok else 'FAIL'}")
band.close(); one.close()
dist.destroy_process_group()
