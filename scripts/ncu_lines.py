"""Per-source-line stall samples of one kernel from an ncu report (needs --import-source on and -lineinfo):
   python scripts/ncu_lines.py rep.ncu-rep <kernel regex> [launch index] [top n]
Prints the CUDA source lines (file:line) with the most warp-stall samples and their dominant stall reasons."""
import csv
import io
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else '0'
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '--kernel-name', 'regex:' + rx,
                      '--launch-skip', skip, '--launch-count', '1'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL).stdout.decode()
rows = list(csv.reader(io.StringIO(out)))
fname, hdr, ix = '', None, None
lines, tot = [], 0
for r in rows:
    if len(r) == 2 and r[0] == 'File Path':
        fname = r[1].split('/')[-1]
        continue
    if len(r) > 10 and r[0] == 'Line No':
        hdr = r
        ix = {}
        for i, h in enumerate(hdr):
            ix.setdefault(h, i)
        continue
    if hdr is None or len(r) < 10 or not r[0]:
        continue            # SASS rows have an empty line number
    try:
        n = int(r[ix['Warp Stall Sampling (All Samples)']] or 0)
    except ValueError:
        continue
    lines.append((n, fname, r))
    tot += n
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
print('total samples', tot)
lines.sort(key=lambda t: -t[0])
for n, f, r in lines[:top]:
    st = sorted(((int(r[ix[s]] or 0), s) for s in stalls), reverse=True)[:3]
    print('%6d %5.1f%%  %-20s %-100s %s inst=%s' % (n, 100.0 * n / max(tot, 1), '%s:%s' % (f, r[0]), r[1].strip()[:100],
                                                  ' '.join('%s:%d' % (s[6:], v) for v, s in st if v), r[ix['Instructions Executed']]))
