"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total
time and share.  usage: summarize_launches.py launches.csv [skip_first_n_launches | last]
`last` = the last whole denoiser eval in the list (from the last single_combine_kernel launch to the end)."""
import csv, sys, collections, re
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith('==')]
rd = csv.DictReader(lines)
for r in rd:
    if r.get('Metric Name') == 'gpu__time_duration.sum':
        v = float(r['Metric Value'].replace(',', ''))
        u = r.get('Metric Unit', 'ns')
        v *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'msecond': 1e3, 'usecond': 1.0, 'nsecond': 1e-3, 'second': 1e6}.get(u, 1e-3)
        rows.append((re.sub(r'\(.*', '', r['Kernel Name']), v))
if len(sys.argv) > 2 and sys.argv[2] == 'last':
    skip = max(i for i, (k, _) in enumerate(rows) if k.startswith('single_combine_kernel'))
else:
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = rows[skip:]
agg = collections.OrderedDict()
for k, v in rows:
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print('launches %d  total %.1f us' % (len(rows), tot))
print('%-50s %6s %12s %10s %7s' % ('kernel', 'count', 'total_us', 'avg_us', 'share'))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print('%-50s %6d %12.1f %10.1f %6.1f%%' % (k[:50], a[0], a[1], a[1] / a[0], 100 * a[1] / tot))
