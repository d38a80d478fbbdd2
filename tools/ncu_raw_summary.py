"""Per-kernel summary of an `ncu --page raw --csv` export (one line per distinct kernel/grid)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
def g(r, k):
    try: return float(r[ix[k]].replace(',', ''))
    except Exception: return float('nan')
def scaled(r, k, to):
    v = g(r, k); u = units[ix[k]]
    f = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'ns': 1e-9, 'us': 1e-6, 'ms': 1e-3, 's': 1}.get(u, 1)
    return v * f / to
def kname(s):          # "void k_hflx<1>(Geo, HflxArgs)" -> "k_hflx" (the name gmc_profile reports)
    s = s.split('(')[0].split('<')[0]
    return s[5:] if s.startswith('void ') else s
seen = set()
print(f"{'kernel':16s} {'grid':>14s} {'ms':>7s} {'rdGB':>6s} {'wrGB':>6s} {'GB/s':>6s} {'sm%':>5s} {'l1%':>5s} {'l2%':>5s} {'fp64%':>5s} {'alu%':>5s} {'lsu%':>5s} {'occ%':>5s} {'regs':>4s} {'inst/pt':>7s} {'ipc':>5s}")
for r in data:
    name = kname(r[ix['Kernel Name']])
    key = (name, r[ix['Grid Size']])
    if key in seen: continue
    seen.add(key)
    t = scaled(r, 'gpu__time_duration.sum', 1e-3)
    rd = scaled(r, 'dram__bytes_read.sum', 1e9); wr = scaled(r, 'dram__bytes_write.sum', 1e9)
    inst = g(r, 'smsp__inst_executed.sum')
    gs = [int(x) for x in r[ix['Grid Size']].strip('()').split(',')]; bs = [int(x) for x in r[ix['Block Size']].strip('()').split(',')]
    thr = gs[0] * gs[1] * gs[2] * bs[0] * bs[1] * bs[2]
    cyc = g(r, 'sm__cycles_elapsed.max') if 'sm__cycles_elapsed.max' in ix else float('nan')
    print(f"{name:16s} {r[ix['Grid Size']].replace(' ', ''):>14s} {t:7.3f} {rd:6.2f} {wr:6.2f} {(rd + wr) / t * 1e3:6.0f} "
          f"{g(r, 'sm__throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} {g(r, 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} "
          f"{g(r, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} {g(r, 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'):5.1f} "
          f"{g(r, 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active'):5.1f} {g(r, 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active'):5.1f} "
          f"{g(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):5.1f} {g(r, 'launch__registers_per_thread'):4.0f} {inst * 32 / thr:7.0f} {inst / (148 * 4) / cyc if cyc == cyc else float('nan'):5.2f}")

if "--json" in sys.argv:
    import json
    out = {"workload": sys.argv[sys.argv.index("--json") + 1], "source": "ncu --set full --clock-control none, one launch per kernel (mean over the captured launches)", "kernels": {}}
    acc = {}
    for r in data:
        name = kname(r[ix['Kernel Name']])
        b = scaled(r, 'dram__bytes_read.sum', 1) + scaled(r, 'dram__bytes_write.sum', 1)
        t = scaled(r, 'gpu__time_duration.sum', 1e-3)
        acc.setdefault(name, []).append((b, t))
    for k, v in acc.items():
        big = [x for x in v if x[0] >= 0.5 * max(y[0] for y in v)]     # ignore the 2-D / few-row launches of the same kernel
        out["kernels"][k] = {"dram_bytes_per_launch": sum(x[0] for x in big) / len(big), "ncu_ms_per_launch": sum(x[1] for x in big) / len(big), "launches": len(big)}
    json.dump(out, open(sys.argv[sys.argv.index("--json") + 2], "w"), indent=1)
