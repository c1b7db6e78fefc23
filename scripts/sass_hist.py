"""Opcode histogram + stall summary of one kernel from `ncu -i rep --page source --csv --print-source sass`.
usage: sass_hist.py file.csv [top_n]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 18
print(rows[0][1][:90])
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
num = lambda r, k: int(r[idx[k]] or 0)
tot = sum(num(r, 'Instructions Executed') for r in data)
samp = sum(num(r, '# Samples') for r in data)
print('warp instructions %d, samples %d, sass lines %d' % (tot, samp, len(data)))
h, hs = collections.Counter(), collections.Counter()
for r in data:
    t = r[idx['Source']].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    h[op] += num(r, 'Instructions Executed')
    hs[op] += num(r, '# Samples')
for k, v in h.most_common(top):
    print('%-10s %10d %5.1f%%  samples %5.1f%%' % (k, v, 100.0 * v / tot, 100.0 * hs[k] / max(samp, 1)))
st = [c for c in hdr if c.startswith('stall_') and 'Not Issued' not in c]
agg = {c: sum(num(r, c) for r in data) for c in st}
s = sum(agg.values())
print({k: round(100.0 * v / s, 1) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
