"""Summarise an `ncu --page source --csv` export: stall-reason totals and the hottest SASS lines.
Usage: python scripts/ncu_source_summary.py file.csv[.gz] [kernel-index] [top-n]"""
import csv, gzip, io, sys
path = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
f = io.TextIOWrapper(gzip.open(path)) if path.endswith('.gz') else open(path)
rows = list(csv.reader(f))
# split per kernel: a "Kernel Name" row starts a section, followed by a header row
secs, cur = [], None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'name': r[1], 'hdr': None, 'data': []}
        secs.append(cur)
    elif cur is not None and cur['hdr'] is None:
        cur['hdr'] = r
    elif cur is not None:
        cur['data'].append(r)
print(len(secs), 'kernel sections')
s = secs[which]
hdr, data = s['hdr'], [r for r in s['data'] if len(r) >= len(s['hdr'])]
idx = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
I = lambda r, h: int(r[idx[h]] or 0)
total = sum(I(r, '# Samples') for r in data)
print(s['name'][:100]); print('total samples', total, 'instructions', len(data))
for st, v in sorted(((st, sum(I(r, st) for r in data)) for st in stalls), key=lambda kv: -kv[1])[:10]:
    print(f'  {st:28s} {v:8d} {v / max(total, 1):.3f}')
for r in sorted(data, key=lambda r: -I(r, '# Samples'))[:topn]:
    dom = max(stalls, key=lambda st: I(r, st))
    print(f"{I(r, '# Samples'):6d} {I(r, 'Instructions Executed'):9d} {dom:22s} {r[idx['Source']].strip()[:100]}")
