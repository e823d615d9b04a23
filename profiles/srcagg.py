"""Aggregate `ncu --page source --csv --print-source cuda,sass` per source line: share of executed instructions and stall samples."""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = defaultdict(lambda: [0, 0, ''])
cur, hdr, seen = None, None, set()
order = []
for r in rows:
    if len(r) == 2 and r[0] == 'File Path':
        cur = r[1].split('/')[-1]
        if cur in seen:          # second profiled launch: stop
            break
        order.append(cur)
        hdr = None
        continue
    if r and r[0] == 'Line No':
        hdr = r
        iI, iS = hdr.index('Instructions Executed'), hdr.index('# Samples')
        continue
    if len(r) == 2 and r[0] == 'Function Name':
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    if r[2] != '':
        try:
            agg[(cur, ln)][0] += int(r[iI]); agg[(cur, ln)][1] += int(r[iS])
        except ValueError:
            pass
    else:
        agg[(cur, ln)][2] = r[1].strip()[:120]
    seen.add(cur) if False else None
tot = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print('total warp-inst', tot, 'samples', ts)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0]}:{k[1]:4d} inst {100*v[0]/tot:5.1f}% samp {100*v[1]/max(ts,1):5.1f}%  {v[2]}")
print('--- by stall samples')
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:15]:
    print(f"{k[0]}:{k[1]:4d} inst {100*v[0]/tot:5.1f}% samp {100*v[1]/max(ts,1):5.1f}%  {v[2]}")
