"""Per-kernel SASS opcode histogram of libwavernn_b200.so (cuobjdump -sass), for profiles/.

  python scripts/sass_histogram.py > profiles/rNN_sass_opcodes.txt
The mnemonics that prove the Blackwell paths (B200_PROFILING.md): UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st,
UBLKCP / UTMALDG = TMA bulk copies, SYNCS = mbarrier; HMMA would be the legacy mma.sync path (none here).
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, 'wavernn_b200', 'libwavernn_b200.so')
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True, check=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r'Function : (\S+)', line)
    if m:
        kern = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip().split('(')[0]
        hist[kern] = collections.Counter()
        continue
    m = re.match(r'\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)', line)
    if m and kern:
        hist[kern][m.group(1).split('.')[0]] += 1
KEY = ['UTCHMMA', 'UTCQMMA', 'UTCBAR', 'UTCATOMSWS', 'LDTM', 'STTM', 'UBLKCP', 'UTMALDG', 'SYNCS', 'LDGSTS', 'HMMA', 'FFMA', 'LDS', 'SHFL',
       'MUFU', 'BAR', 'LDG', 'STG', 'RED', 'ATOM']
print('SASS opcode histogram per kernel, ' + os.path.relpath(lib, ROOT) + ' (sm_100a)')
print('%-44s %7s  ' % ('kernel', 'instr') + ' '.join('%7s' % k[:7] for k in KEY))
tot = collections.Counter()
for k, c in hist.items():
    n = sum(c.values())
    if n == 0:
        continue
    tot.update(c)
    print('%-44s %7d  ' % (k[-44:], n) + ' '.join('%7d' % c.get(x, 0) for x in KEY))
print('%-44s %7d  ' % ('ALL KERNELS', sum(tot.values())) + ' '.join('%7d' % tot.get(x, 0) for x in KEY))
print('\ntensor-core path: UTCHMMA (tcgen05.mma kind::f16) %d, LDTM (tcgen05.ld) %d, UBLKCP (cp.async.bulk) %d, HMMA (legacy mma.sync) %d'
      % (tot['UTCHMMA'], tot['LDTM'], tot['UBLKCP'], tot['HMMA']))
