"""Summarise an `ncu --page source --csv --print-source sass` dump: opcode histogram and the
instructions with the most stall samples.  usage: ncu_src.py <csv> [kernel index] [top n]"""
import csv, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
ki = int(sys.argv[2]) if len(sys.argv) > 2 else 0
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
kern, cur = [], None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'name': r[1], 'hdr': None, 'rows': []}; kern.append(cur); continue
    if cur is None: continue
    if cur['hdr'] is None: cur['hdr'] = r; continue
    cur['rows'].append(r)
k = kern[ki]; h = k['hdr']
ia, isamp, isrc = h.index('Instructions Executed'), h.index('# Samples'), h.index('Source')
st = [i for i, x in enumerate(h) if x.startswith('stall_') and 'Not Issued' not in x]
tot = sum(int(r[ia]) for r in k['rows']); ts = sum(int(r[isamp]) for r in k['rows'])
print(k['name'], 'warp instr', tot, 'samples', ts)
c, cs = Counter(), Counter()
for r in k['rows']:
    t = r[isrc].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    c[op] += int(r[ia]); cs[op] += int(r[isamp])
for op, n in c.most_common(18): print(f'{op:10s} {n:12d} {100*n/tot:5.1f}%  samples {100*cs[op]/ts:5.1f}%')
agg = Counter()
for r in k['rows']:
    for j in st: agg[h[j][6:]] += int(r[j])
print({a: round(100*b/ts, 1) for a, b in agg.most_common(8)})
top = sorted(range(len(k['rows'])), key=lambda i: -int(k['rows'][i][isamp]))[:topn]
for i in sorted(top):
    r = k['rows'][i]
    s = {h[j][6:]: int(r[j]) for j in st if int(r[j]) > 0}
    s = dict(sorted(s.items(), key=lambda kv: -kv[1])[:3])
    print(i, r[isrc][:64].ljust(64), r[ia], round(100*int(r[isamp])/ts, 2), s)
