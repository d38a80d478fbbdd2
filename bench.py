#!/usr/bin/env python
"""bench.py — simulated days per wall-clock day of the GMCORE dynamics hot path on B200.

Contract: `python bench.py --gpus N --steps K --warmup W [--impl reference]` prints ONE
JSON line (rank 0).  A "step" is one model time step (time_integrate + time_advance +
diagnose, gmcore_mod.F90:138-144) of the named workload on synthetic (analytic) input.

  value     device-resident throughput (state already in HBM), CUDA events on the
            library's stream, max over ranks
  e2e       the same job through the C ABI from HOST buffers: gmc_upload_state from
            pinned memory + gmc_prepare_run + K x (gmc_time_integrate + gmc_diagnose,
            a 64-byte D2H read per step) + gmc_download_state, wall clock around it
  roofline  dominant kernel: algorithmic bytes per launch / mean launch duration
            (per-launch CUDA events inside the library, gmc_profile) vs the measured
            HBM copy bandwidth in MEASURED_PEAKS.json
  cpu_baseline  the CPU restatement (oracle/, "port") timed on the host cores on a
            bounded sample of the same workload
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # Williamson 5 at 0.5 deg, shallow water
    "swm_mz_720x360": dict(nlon=720, nlat=360, nlev=1, dt=75.0, case="mountain_zonal_flow", bytes_per_cell=936),
    # Jablonowski-Williamson baroclinic wave, hydrostatic, 1 FFSL tracer (qv)
    "bw_360x180x32": dict(nlon=360, nlat=180, nlev=32, dt=150.0, case="baroclinic_wave", bytes_per_cell=2320),
    "bw_1440x720x32": dict(nlon=1440, nlat=720, nlev=32, dt=30.0, case="baroclinic_wave", bytes_per_cell=2320),
}
# BASELINE.json quotes its metric on the 0.25 deg x 32-level baroclinic wave; it fits one B200
# (about 27 GB of device state), so that is the N=1 workload.
DEFAULT_WORKLOAD = "bw_1440x720x32"
# Fallback CPU sample ONLY when the host cannot hold the oracle state (~80 GB at 0.25 deg): same case and
# levels on a coarser grid, cost scaled by the cell count, and labelled as such in the line.
CPU_SAMPLE = {"bw_1440x720x32": "bw_360x180x32"}

# Algorithmic 3-D r8 array passes per launch of each kernel (SURVEY.md §8d / DESIGN.md):
# reads + writes of full 3-D fields; 2-D fields and coefficient vectors ignored.
KERNEL_PASSES = {
    "k_fwd1": 3 + 7, "k_pv_edge": 5 + 2, "k_tend": 5 + 3, "k_uv_update": 5 + 2,
    "k_filter_rows": 7.0 / 3,      # mean of the three shallow-water launches: gz -> gz_f, gz, m (1+3); u -> u_f (1+1); v -> v_f (1+1)
    "k_diag_partial": 8,
}
# hydrostatic core (k_tend / k_uv_update read more there); k_hflx: 5 reads on the inner call
# (mx == my) and 6 on the outer call + 2 writes -> mean 7.5
KERNEL_PASSES_HYDRO = {
    "k_hflx": 7.5, "k_ffsl_march_in": 7 + 2, "k_ffsl_march_out": 6 + 2, "k_ffsl_x_in": 4 + 1, "k_ffsl_y_in": 4 + 1,
    "k_ffsl_x_out": 3 + 1, "k_ffsl_y_out": 3 + 1, "k_ffsl_xw_in": 4 + 1, "k_ffsl_xw_out": 3 + 1, "k_fwd1": 3 + 7, "k_tend": 9 + 2, "k_cfl": 2 + 4, "k_pt_col": 7 + 1, "k_we_col": 3 + 2,
    "k_uv_update_h": 6 + 2, "k_gz_col": 1 + 2, "k_tracer_v": 4 + 1, "k_ffsl_q": 5 + 2, "k_pv_edge": 5 + 2,
    "k_ph": 0 + 3, "k_accum": 2 + 2, "k_diag_partial": 10, "k_filter_rows": 1 + 1,
}
L2_BYTES = 126e6


def ncu_traffic(kernel, workload):
    """DRAM bytes per launch of `kernel` from the committed `ncu --set full` capture of this workload
    (profiles/r02_ncu_traffic.json, written by tools/ncu_raw_summary.py --json), or None."""
    p = os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")
    try:
        t = json.load(open(p))
        return t["kernels"][kernel]["dram_bytes_per_launch"] if t.get("workload") == workload else None
    except Exception:
        return None


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.device)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def mark(self):
        """Samples before this point (warm-up, idle waiting) are dropped if later ones exist."""
        self.n0 = len(self.rows)

    def wait_first(self, timeout=5.0):
        t0 = time.time()
        while self.proc and not self.rows and time.time() - t0 < timeout:
            time.sleep(0.02)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        rows = self.rows[getattr(self, 'n0', 0):] or self.rows
        self.rows = rows
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 9:
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def max_over_ranks(x, world):
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def pinned_like(a):
    import torch
    t = torch.empty(a.shape, dtype=torch.float64, pin_memory=True)
    n = t.numpy(); n[...] = a
    return n, t


def bench_tracer(host):
    """The advected tracer of the timed workloads: a positive, smooth specific-humidity-like field (wavenumber-2 in
    longitude, moist tropics, decaying upwards) so that the FFSL limiter, the negative-filling sweeps and lon_damp see
    real gradients (an all-zero field would be a best case)."""
    q = np.zeros((host.cfg.ntracers,) + host.shape("cell"))
    lon = host.vec["full_lon"][host.hx:host.hx + host.nlon]
    lat = host.vec["full_lat"][host.hy:host.hy + host.nlat]
    for l in range(host.cfg.ntracers):
        f = 1.0e-2 * (np.cos(lat)[:, None] ** 2 * (1.0 + 0.5 * np.sin((l + 2) * lon)[None, :]) + 0.05)
        for k in range(host.nlev):
            host.interior(q[l], "cell")[k] = f * ((k + 1.0) / host.nlev) ** 2
    return q


def oracle_mem_gb(wl):
    """Host memory of one oracle instance: ~225 halo'd r8 arrays (the reference's own allocation)."""
    hx = 130 if wl["nlon"] >= 1440 else 90
    return 225 * (wl["nlon"] + 2 * hx) * (wl["nlat"] + 4) * (wl["nlev"] + 4) * 8 / 1e9 + 1.0


def mem_available_gb():
    try:
        return [int(l.split()[1]) for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0] / 1e6
    except Exception:
        return 8.0


def cpu_baseline(wl, name, budget_s=20.0, min_steps=2, max_steps=200):
    """Time the CPU restatement (oracle/, the -O3 OpenMP build liboracle_fast.so) on the host cores.

    ONE instance of the named workload itself on ALL host cores: the level loops of the restatement are OpenMP-
    parallel (oracle/orc_core.hpp ORC_PAR), the shared-memory stand-in for the reference's MPI decomposition.  The
    sample is bounded in STEPS (1 warm-up + >= min_steps timed), never in grid size, unless the host cannot hold the
    ~80 GB oracle state of the 0.25 degree case: then the same case on the coarser grid of CPU_SAMPLE is timed and the
    line says so ("workload_timed" / "scaled_to")."""
    cores = os.cpu_count() or 1
    timed_name, twl, scale = name, wl, 1.0
    if oracle_mem_gb(wl) > 0.8 * mem_available_gb() and name in CPU_SAMPLE:
        timed_name = CPU_SAMPLE[name]; twl = WORKLOADS[timed_name]
        scale = (wl["nlon"] * wl["nlat"] * wl["nlev"]) / (twl["nlon"] * twl["nlat"] * twl["nlev"])
    env = dict(os.environ, OMP_NUM_THREADS=str(cores), OMP_PROC_BIND="close", OMP_PLACES="cores")
    t0 = time.perf_counter()
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cpu_worker.py"), timed_name, str(budget_s), str(min_steps), str(max_steps)],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env)
    wall = time.perf_counter() - t0
    try:
        n, el, rss, threads = p.stdout.split()[-4:]
        n, el, rss, threads = int(n), float(el), float(rss), int(threads)
    except Exception:
        raise RuntimeError("CPU baseline worker failed: " + p.stderr[-300:])
    s_per_step = el / n * scale
    sample = (f"1 instance of {timed_name} on {threads} OpenMP threads (liboracle_fast.so: g++ -O3 -march=x86-64-v3 -fopenmp, "
              f"level loops parallel): {n} steps timed after 1 warm-up step, {el / n:.2f} s/step, rss {rss / 1024:.1f} GB, "
              f"{wall:.0f} s wall incl. set-up; host has {cores} cores")
    if scale != 1.0:
        sample += f"; host memory too small for {name} ({oracle_mem_gb(wl):.0f} GB): cost per step scaled x{scale:g} by cell count"
    return {"value": wl["dt"] / s_per_step, "unit": "simulated days/day", "cores": threads, "kind": "port", "sample": sample,
            "s_per_step": s_per_step, "workload_timed": timed_name, "steps_timed": n}


def run_reference(args, name, wl, rank, world, out=sys.stdout):
    """--impl reference: the reference's CPU implementation of the path.  The Fortran itself cannot be built (no
    Fortran compiler, MPI or netCDF on this image or on the GPU box -- probed), so this is the
    oracle port on all host cores, on this arm's own config.  Rank 0 only."""
    if rank != 0:
        return
    t0 = time.perf_counter()
    # steps: the requested K, bounded so that the whole run ends within a few minutes (>= 3 timed steps)
    cb = cpu_baseline(wl, name, budget_s=150.0, min_steps=min(3, max(1, args.steps)), max_steps=max(1, args.steps))
    cfg = {"workload": name, **{k: wl[k] for k in ("nlon", "nlat", "nlev", "dt", "case")}, "time_scheme": "wrfrk3"}
    if cb["workload_timed"] != name:
        cfg["workload_timed"] = cb["workload_timed"]; cfg["scaled_to"] = name
    line = {"impl": "reference", "metric": "simulated days per wall-clock day", "value": cb["value"], "unit": "simulated days/day",
            "n_gpus": args.gpus, "steps": cb["steps_timed"], "steps_requested": args.steps, "warmup": 1, "ms_per_step": cb["s_per_step"] * 1e3,
            "higher_is_better": True, "scaling": "strong" if wl["nlev"] > 1 else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "simulated days/day", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), file=out, flush=True)


def main():
    # stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner) go to stderr
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    try:
        _main(real_stdout)
    finally:
        real_stdout.flush()


def _main(out):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default=os.environ.get("GMCORE_WORKLOAD", DEFAULT_WORKLOAD))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    name = args.workload; wl = WORKLOADS[name]
    K, W = args.steps, max(3, args.warmup)

    if args.impl == "reference":
        rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
        run_reference(args, name, wl, rank, world, out)
        return

    rank, world, local = dist_setup(args.gpus)
    import torch
    from gmcore_b200.host import Host
    from gmcore_b200.model import GmcoreGPU

    host = Host(wl["nlon"], wl["nlat"], wl["nlev"], wl["dt"], wl["case"])
    ic = host.initial_condition()
    banded = world > 1 and wl["nlev"] > 1      # latitude bands + NCCL halo exchange (strong scaling)
    if banded:
        from gmcore_b200.model import make_band_nccl
        gpu = make_band_nccl(host, rank, world, local)
    else:
        gpu = GmcoreGPU(host, device=local)
    # N > 1: the serial-vs-parallel check of the NCCL path, outside every timed region (small grid, this rank's band
    # over ncclSend/Recv with the overlap on against one band on the same device; bar 1e-13)
    bpar = None
    if banded:
        from gmcore_b200.model import band_parity
        bpar = band_parity(rank, world, local)
    gpu.set_state(ic)
    q0 = None
    if host.cfg.ntracers:
        q0 = bench_tracer(host)
        gpu.set_tracers(q0)
    gpu.prepare_run()
    # nvidia-smi takes a while to deliver its first sample (longer with 8 ranks starting one each): start it before
    # the warm-up steps (same load) and make sure it is delivering before the timed region begins
    clocks = ClockSampler(local); clocks.start()
    gpu.run_steps(W, diag=True)
    gpu.synchronize()
    clocks.wait_first()

    n3 = wl["nlon"] * wl["nlat"] * wl["nlev"]
    # cells this rank's kernels process (its own rows; the redundant halo rows are not counted)
    n3_k = wl["nlon"] * (gpu.cfg.lat_iend - gpu.cfg.lat_ibeg + 1) * wl["nlev"] if banded else n3
    working_set = 40 * n3_k * 8
    flush = working_set < 2 * L2_BYTES
    flush_buf = torch.empty(int(2 * L2_BYTES) // 8, dtype=torch.float64, device=f"cuda:{local}") if flush else None

    # ---- device-resident timed region: exactly K steps --------------------------------
    barrier(world); torch.cuda.synchronize()
    clocks.mark()
    l0 = gpu.launches
    if flush:
        ms = 0.0
        for _ in range(K):
            flush_buf.fill_(1.0); torch.cuda.synchronize()
            gpu.timer_start(); gpu.run_steps(1, diag=True); ms += gpu.timer_stop_ms()
    else:
        gpu.timer_start(); gpu.run_steps(K, diag=True); ms = gpu.timer_stop_ms()
    torch.cuda.synchronize(); barrier(world)
    launches = gpu.launches - l0
    ms = max_over_ranks(ms, world)
    # bands: the N ranks integrate ONE domain (strong scaling); shallow water: N replicas
    value = (1 if banded else world) * K * wl["dt"] / (ms / 1e3)

    # ---- per-kernel CUDA-event pass (same K steps, profiling on) -------------------------
    gpu.profile(True); gpu.run_steps(K, diag=True); rep = gpu.profile_report(); gpu.profile(False)
    clk = clocks.stop()      # sampled over the timed region and this identical-load profiling pass (20 ms period)
    tot = sum(v[1] for v in rep.values())
    top = max(rep.items(), key=lambda kv: kv[1][1])
    peak, peak_src = measured_peak_gbs()
    passes = (KERNEL_PASSES_HYDRO if wl["nlev"] > 1 else KERNEL_PASSES).get(top[0])
    mean_ms = top[1][1] / top[1][0]
    achieved = (passes * 8 * n3_k / (mean_ms * 1e-3) / 1e9) if passes else None
    step_gbs = wl["bytes_per_cell"] * n3 / (ms / K * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": top[0], "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None,
                "traffic": ncu_traffic(top[0], name) if world == 1 else None,
                "algorithmic_bytes_per_launch": (passes * 8 * n3_k) if passes else None,
                "kernel_share_of_step": top[1][1] / tot, "kernel_mean_ms": mean_ms,
                "whole_step_algorithmic_gbs": step_gbs, "whole_step_frac": step_gbs / (peak * (world if banded else 1)),
                "kernels_ms_per_step": {k: v[1] / K for k, v in sorted(rep.items(), key=lambda kv: -kv[1][1])}}

    # ---- end to end through the C ABI from pinned host buffers ------------------------------
    gpu2 = make_band_nccl(host, rank, world, local) if banded else GmcoreGPU(host, device=local)
    # this rank's host arrays in pinned memory (banded: its own rows -- slicing the global arrays is host
    # set-up, a banded Fortran host owns only its rows in the first place)
    src = gpu2.slice_state(ic) if banded else ic
    icp = {k: (pinned_like(v)[0] if v is not None else None) for k, v in src.items()}
    q0p = None
    if q0 is not None:
        q0p = pinned_like(gpu2.slice_tracers(q0) if banded else q0)[0]
    # page-locked destination of the final download: the host model's own state arrays, allocated once
    outp = {k: (pinned_like(np.zeros_like(v))[0] if v is not None else None) for k, v in src.items()}
    # bytes that cross PCIe: the library copies the interior of each halo'd host array (one strided copy per field)

    def wire_bytes(a):
        if a is None:
            return 0
        rows = a.shape[-2] - 2 * host.hy                      # owned rows of this array (full- or half-lat)
        levs = 1 if a.ndim == 2 else (a.shape[0] - 4 if a.shape[0] == wl["nlev"] + 4 else a.shape[0])
        return 8 * wl["nlon"] * rows * levs
    h2d = sum(wire_bytes(v) for v in icp.values())
    # untimed warm-up of this handle (first-use costs: NCCL connection set-up, pinned-page registration)
    (gpu2.set_state_sliced if banded else gpu2.set_state)(icp)
    if q0 is not None:
        (gpu2.set_tracers_sliced if banded else gpu2.set_tracers)(q0p)
    gpu2.prepare_run(); gpu2.time_integrate(); gpu2.diagnose(); gpu2.synchronize()
    gpu2.reset_clock()
    barrier(world); torch.cuda.synchronize()
    t0 = time.perf_counter()
    (gpu2.set_state_sliced if banded else gpu2.set_state)(icp)
    if q0 is not None:
        (gpu2.set_tracers_sliced if banded else gpu2.set_tracers)(q0p); h2d += sum(wire_bytes(x) for x in q0p)
    t1 = time.perf_counter()
    gpu2.prepare_run()
    for _ in range(K):
        gpu2.time_integrate()
        gpu2.diagnose()
    t2 = time.perf_counter()
    st = gpu2.download_state(out=outp)
    gpu2.synchronize()
    t3 = time.perf_counter()
    e2e_s = max_over_ranks(t3 - t0, world)
    # host timestamps of rank 0 at the phase boundaries (no synchronisation added: the uploads, every diagnose and the
    # download are synchronous calls already)
    phases = {"upload": (t1 - t0) * 1e3, "prepare_run+steps": (t2 - t1) * 1e3, "download": (t3 - t2) * 1e3}
    d2h = sum(wire_bytes(v) for v in st.values())
    e2e = {"value": (1 if banded else world) * K * wl["dt"] / e2e_s, "unit": "simulated days/day",
           "h2d_bytes_per_step": h2d / K, "d2h_bytes_per_step": d2h / K + 64, "bytes_are": "per rank", "phases_ms_rank0": phases,
           "region": "upload_state(pinned)+prepare_run+K*(time_integrate+diagnose D2H)+download_state"}
    gpu2.close()

    if banded:
        from gmcore_b200.model import balanced_rows, round_robin
        band_rows = os.environ.get("GMCORE_BAND_ROWS", "round_robin")
        parts = balanced_rows(host, world) if band_rows == "balanced" else round_robin(host.nlat, world)
        band_sizes = "/".join(str(e - b + 1) for b, e in parts)
    if rank == 0:
        cb = None
        if not args.no_cpu_baseline:
            try:
                cb = cpu_baseline(wl, name, budget_s=25.0, min_steps=2, max_steps=50)
                for k in ("s_per_step", "workload_timed", "steps_timed"):
                    cb.pop(k)
            except Exception as e:      # never lose the GPU line to a CPU-side failure
                cb = {"value": None, "unit": "simulated days/day", "cores": 0, "kind": "port", "sample": f"failed: {e}"}
        line = {"metric": "simulated days per wall-clock day", "value": value, "unit": "simulated days/day", "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
                "scaling": "strong" if wl["nlev"] > 1 else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": name, **{k: wl[k] for k in ("nlon", "nlat", "nlev", "dt", "case")},
                           "time_scheme": "wrfrk3", "hx": host.hx, "parallelism": (f"{world} latitude bands ({band_rows} rows: {band_sizes}), NCCL halo exchange" if banded else f"replicas x{world}") if world > 1 else "1 GPU",
                           "l2": "flushed between timed steps (256 MB write)" if flush else "working set >> L2 (no flush)"},
                "gpu_launches": launches, "clocks": clk, "e2e": e2e, "roofline": roofline, "cpu_baseline": cb}
        if bpar is not None:
            line["band_parity"] = bpar
        print(json.dumps(line), file=out, flush=True)
    gpu.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
