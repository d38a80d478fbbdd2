"""Host-side logic of the latitude-band decomposition (process_mod.F90:93-139, parallel_types_mod.F90:235-262)
exercised with two real processes over gloo: partition, band slicing, the S/N halo rows a neighbour
must supply (fill_halo, parallel_mod.F90:501-527) and re-assembly of the global field."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gmcore_b200.host import Host
from gmcore_b200.model import band_config, band_slice, round_robin, scatter_rows


def test_round_robin_matches_reference_rule():
    # remainder rows go to the lowest ranks; contiguous, covering 1..n
    for n, p in ((720, 8), (181, 4), (90, 7), (33, 2)):
        parts = round_robin(n, p)
        assert parts[0][0] == 1 and parts[-1][1] == n
        sizes = [e - b + 1 for b, e in parts]
        assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
        assert all(parts[r + 1][0] == parts[r][1] + 1 for r in range(p - 1))


def test_balanced_rows_cover_the_sphere_and_relieve_the_polar_bands():
    from gmcore_b200.model import balanced_rows, row_weights
    for (nlon, nlat, nlev, dt), ps in (((1440, 720, 32, 30.0), (1, 2, 4, 8)), ((120, 61, 8, 300.0), (2, 3, 5, 8)), ((48, 25, 4, 600.0), (2, 3, 4))):
        host = Host(nlon, nlat, nlev, dt, "baroclinic_wave")
        w = row_weights(host)
        # half row j is owned with full row j, so the two hemispheres mirror each other up to that one-row shift
        assert w.shape == (nlat,) and w.min() >= 1.0 and abs(w[:nlat // 2].sum() / w[(nlat + 1) // 2:].sum() - 1) < 0.02
        assert w[1] > w[nlat // 2]                                                    # polar rows carry the wide filters
        for p in ps:
            parts = balanced_rows(host, p)
            assert len(parts) == p and parts[0][0] == 1 and parts[-1][1] == nlat
            assert all(parts[r + 1][0] == parts[r][1] + 1 for r in range(p - 1))
            sizes = [e - b + 1 for b, e in parts]
            assert min(sizes) >= (6 if p > 1 else nlat)                               # gmc_create's minimum band
            if p >= 3:
                assert sizes[0] <= min(sizes[1:-1]) and sizes[-1] <= min(sizes[1:-1])
    host = Host(1440, 720, 32, 30.0, "baroclinic_wave")
    cost = lambda parts: max(row_weights(host)[b - 1:e].sum() for b, e in parts)
    assert cost(balanced_rows(host, 8)) < cost(round_robin(720, 8))
    assert band_config(host, 0, 8, 0, rows="balanced").lat_iend == balanced_rows(host, 8)[0][1]


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        host = Host(48, 25, 4, 600.0, "baroclinic_wave")
        hy = host.hy
        ic = host.initial_condition()
        cfg = band_config(host, rank, world, 0)
        ok = True
        for name, loc in (("pt_or_gz", "cell"), ("v_lat", "lat"), ("phs", "2d")):
            glob = ic[name]
            band = band_slice(host, cfg, glob, loc)
            nrow = band.shape[-2]
            own = band[..., hy:nrow - hy, :]
            # exchange: my first/last hy owned rows go to the south/north neighbour
            recv_s = torch.zeros_like(torch.from_numpy(np.ascontiguousarray(own[..., :hy, :])))
            recv_n = torch.zeros_like(recv_s)
            reqs = []
            if rank > 0:
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(own[..., :hy, :])), rank - 1))
                reqs.append(dist.irecv(recv_s, rank - 1))
            if rank + 1 < world:
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(own[..., -hy:, :])), rank + 1))
                reqs.append(dist.irecv(recv_n, rank + 1))
            for r in reqs:
                r.wait()
            # the rows received are exactly the halo rows of the band slice of the global array
            if rank > 0:
                ok &= np.array_equal(recv_s.numpy(), band[..., :hy, :])
            if rank + 1 < world:
                ok &= np.array_equal(recv_n.numpy(), band[..., nrow - hy:, :])
            # re-assembly: every rank scatters its owned rows, the sum over ranks is the global interior
            part = np.zeros_like(glob)
            scatter_rows(host, cfg, band, part, loc)
            t = torch.from_numpy(part)
            dist.all_reduce(t)
            interior = host.interior(t.numpy(), loc); ref = host.interior(glob, loc)
            ok &= np.array_equal(interior, ref)
        # global sum (diagnose): per-band partial sums all-reduced == single-process sum to round-off
        m = band_slice(host, cfg, ic["pt_or_gz"], "cell")
        s = torch.tensor([m[..., hy:m.shape[-2] - hy, host.hx:-host.hx][2:-2].sum()], dtype=torch.float64)
        dist.all_reduce(s)
        tot = host.interior(ic["pt_or_gz"], "cell").sum()
        ok &= abs(s.item() - tot) <= 1e-13 * abs(tot)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_rank_band_exchange_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=100) for _ in procs)
    for p in procs:
        p.join(30)
    assert res == {0: True, 1: True}
