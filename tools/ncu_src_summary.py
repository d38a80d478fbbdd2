"""Summarise an `ncu --page source --csv` export: stall mix, dynamic opcode mix, hottest SASS lines."""
import csv, collections, sys
for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
    st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    tot = collections.Counter(); ins = 0; op = collections.Counter(); hot = []
    for r in rows[2:]:
        if len(r) < len(hdr) or not r[0].startswith('0x'): continue
        for s in st:
            try: tot[s] += int(r[ix[s]])
            except ValueError: pass
        try: n = int(r[ix['Instructions Executed']])
        except ValueError: n = 0
        ins += n
        toks = [x for x in r[ix['Source']].split() if not x.startswith('@')]
        op[toks[0].split('.')[0]] += n
        try: hot.append((int(r[ix['# Samples']]), r[ix['Source']].strip()))
        except ValueError: pass
    T = sum(tot.values()) or 1
    print(rows[0][1], 'samples', T, 'warp-inst', ins)
    print('  stalls:', ', '.join(f"{s[6:]}={100*v/T:.0f}%" for s, v in tot.most_common(8)))
    print('  ops:', ', '.join(f"{o}={100*v/max(ins,1):.0f}%" for o, v in op.most_common(16)))
    if '-v' in sys.argv: 
        for s, src in sorted(hot, reverse=True)[:25]: print(f"   {s:6d} {src}")
