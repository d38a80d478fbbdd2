"""Executed warp instructions per CUDA source line: joins an `ncu --page source --csv` export (per-SASS-instruction
counts) with `nvdisasm -g` output of the same kernel (line markers).  Usage: ncu_lines.py src.csv kernel.sass [N]"""
import csv, re, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
cnt = []; stall = []
for r in rows[2:]:
    if len(r) < len(hdr) or not r[0].startswith('0x'): continue
    try: cnt.append(int(r[ix['Instructions Executed']]))
    except ValueError: cnt.append(0)
    try: stall.append(int(r[ix['# Samples']]))
    except ValueError: stall.append(0)
line = None; k = 0
by = collections.Counter(); st = collections.Counter(); ops = collections.defaultdict(collections.Counter)
for l in open(sys.argv[2]):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: line = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s*/\*[0-9a-f]{4,}\*/\s+(.*?);', l)
    if m and k < len(cnt):
        by[line] += cnt[k]; st[line] += stall[k]
        op = [t for t in m.group(1).split() if not t.startswith('@')][0].split('.')[0]
        ops[line][op] += cnt[k]; k += 1
tot = sum(by.values()); ts = sum(st.values()) or 1
print("instructions matched", k, "of", len(cnt), "total executed", tot)
N = int(sys.argv[3]) if len(sys.argv) > 3 else 40
for ln, n in by.most_common(N):
    print(f"{100*n/tot:5.1f}% inst {100*st[ln]/ts:5.1f}% stall  {ln[0]}:{ln[1]}  " + ' '.join(f"{o}:{100*v/tot:.1f}" for o, v in ops[ln].most_common(5)))
