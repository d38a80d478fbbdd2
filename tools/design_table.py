"""DESIGN.md §3.1 table from a tools/kprof.py listing + profiles/r02_ncu_traffic.json.
usage: python tools/design_table.py kprof.txt [peak_gbs]"""
import json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import KERNEL_PASSES_HYDRO, WORKLOADS, measured_peak_gbs
txt = open(sys.argv[1]).read()
name = re.search(r"^(\S+) lib=", txt, re.M).group(1)
wl = WORKLOADS[name]; n3 = wl["nlon"] * wl["nlat"] * wl["nlev"]
peak = float(sys.argv[2]) if len(sys.argv) > 2 else measured_peak_gbs()[0]
t = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")))["kernels"]
print(f"| kernel | launches/step | ms/launch (CUDA events) | alg. passes | alg. GB/s | % of measured HBM peak ({peak:.0f} GB/s) | ncu DRAM GB/launch |")
print("|---|---|---|---|---|---|---|")
tot_ms = 0.0
for m in re.finditer(r"^\s+(k_\w+)\s+n/step=\s*([\d.]+) ms/step=\s*([\d.]+) mean_ms=\s*([\d.]+)", txt, re.M):
    k, n, ms, mean = m.group(1), float(m.group(2)), float(m.group(3)), float(m.group(4)); tot_ms += ms
    if ms < 0.1: continue
    p = KERNEL_PASSES_HYDRO.get(k)
    gbs = p * 8 * n3 / (mean * 1e-3) / 1e9 if p else None
    d = t.get(k, {}).get("dram_bytes_per_launch")
    print(f"| `{k}` | {n:g} | {mean:.3f} | {p if p else '—'} | {gbs:.0f} | {100 * gbs / peak:.0f} % | {d / 1e9:.2f} |" if p else
          f"| `{k}` | {n:g} | {mean:.3f} | — | — | — | {(d or 0) / 1e9:.2f} |")
m = re.search(r"ms/step=([\d.]+) SDPD=([\d.]+)", txt)
print(f"| whole step | | {float(m.group(1)):.2f} ms (kernels sum {tot_ms:.2f}) | 290 | {290 * 8 * n3 / (float(m.group(1)) * 1e-3) / 1e9:.0f} | {100 * 290 * 8 * n3 / (float(m.group(1)) * 1e-3) / 1e9 / peak:.0f} % | |")
