"""Opcode summary of every kernel in librotgeo.so (cuobjdump -sass): the mnemonics that prove the sm_100a
features the design relies on, and local-memory traffic.   python tools/sass_summary.py > profiles/rNN_sass_opcodes.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, 'orientedobjectdetection_b200', 'librotgeo.so')
out = subprocess.run(['cuobjdump', '-sass', so], capture_output=True, text=True).stdout
WATCH = ['UBLKCP', 'UGETNEXTWORKID', 'SYNCS', 'UCGABAR', 'MAPA', 'ACQBULK', 'LDL', 'STL', 'ATOMS', 'ATOMG', 'RED', 'VOTE', 'MATCH',
         'MUFU', 'FFMA', 'DFMA', 'LDS', 'STS', 'LDG', 'STG', 'BAR', 'SHFL', 'ERRBAR', 'CCTL']
kern, counts = None, collections.OrderedDict()
for line in out.split('\n'):
    m = re.search(r'Function : (\S+)', line)
    if m:
        kern = m.group(1)
        counts[kern] = collections.Counter()
        continue
    m = re.search(r'/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)', line)
    if m and kern:
        op = m.group(1)
        counts[kern]['_total'] += 1
        for w in WATCH:
            if op.startswith(w):
                counts[kern][w] += 1
demangle = subprocess.run(['c++filt'] + list(counts), capture_output=True, text=True).stdout.split('\n')
print('# SASS opcode counts per kernel of librotgeo.so (sm_100a), static instruction counts')
print('# UBLKCP = TMA bulk copy, UGETNEXTWORKID = cluster launch control (work stealing), SYNCS = mbarrier ops,')
print('# UCGABAR = cluster barrier, MAPA = distributed-shared-memory address mapping, LDL/STL = local memory')
for (k, c), name in zip(counts.items(), demangle):
    short = re.sub(r'\((?!anonymous).*', '', name.replace('(anonymous namespace)::', ''))[:70]
    items = ' '.join('%s=%d' % (w, c[w]) for w in WATCH if c[w])
    print('%-72s total=%-5d %s' % (short, c['_total'], items))
