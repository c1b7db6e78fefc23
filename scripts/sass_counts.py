"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / bulk-copy use, from `cuobjdump -sass` of the built
library (no GPU needed).  usage: sass_counts.py [lib.so] > profiles/rNN_sass_tcgen05_counts.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'genie2_b200', 'libgenie2_b200.so')
out = subprocess.run(['cuobjdump', '-sass', lib], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL).stdout.decode()
OPS = ['UTCHMMA', 'UTCQMMA', 'UTCBAR', 'LDTM', 'STTM', 'UTCCP', 'UBLKCP', 'UBLKPF', 'UTMALDG', 'UTMASTG', 'SYNCS', 'ELECT', 'MUFU', 'FFMA2', 'HMMA', 'ACQBULK']
cnt = collections.OrderedDict()
cur, arch = None, None
for line in out.splitlines():
    m = re.match(r'\s*Function : (\S+)', line)
    if m:
        cur = subprocess.run(['c++filt', m.group(1)], stdout=subprocess.PIPE).stdout.decode().strip()
        cur = re.sub(r'\(.*', '', cur).replace('void ', '')
        cnt[cur] = collections.Counter()
        continue
    m = re.match(r'\s*arch = (\S+)', line)
    if m:
        arch = m.group(1)
    if cur is None:
        continue
    m = re.match(r'\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', line)
    if m:
        op = m.group(1).split('.')[0]
        cnt[cur]['_total'] += 1
        if op in OPS:
            cnt[cur][op] += 1
print('library %s   arch %s' % (os.path.basename(lib), arch))
print('%-52s %7s ' % ('kernel', 'instr') + ' '.join('%7s' % o for o in OPS))
for k, c in cnt.items():
    if c['UTCHMMA'] or c['LDTM'] or c['UBLKCP']:
        print('%-52s %7d ' % (k[:52], c['_total']) + ' '.join('%7d' % c[o] for o in OPS))
print('\nkernels without tensor-core instructions: ' + ', '.join(k for k, c in cnt.items() if not (c['UTCHMMA'] or c['LDTM'] or c['UBLKCP'])))
