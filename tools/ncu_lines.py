"""Per-source-line view of an `ncu --set full --import-source on` capture of one librotgeo kernel:
executed warp instructions, stall samples, global-memory sectors, aggregated over the SASS
instructions of each source line (the offset -> line table comes from `nvdisasm -g` of the
cubins inside the library; ncu's own source page needs the GUI for this join).

    python tools/ncu_lines.py <capture.ncu-rep> <mangled-name substring> [lines=16]
"""
import csv, glob, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, 'orientedobjectdetection_b200', 'csrc')
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 16

tmp = tempfile.mkdtemp()
subprocess.run(['cuobjdump', '-xelf', 'all', os.path.join(ROOT, 'orientedobjectdetection_b200', 'librotgeo.so')], cwd=tmp,
               check=True, capture_output=True)
line_of = {}
for cubin in glob.glob(os.path.join(tmp, '*.cubin')):
    txt = subprocess.run(['nvdisasm', '-g', '-c', cubin], capture_output=True, text=True).stdout
    if kern not in txt:
        continue
    cur, inside = None, False
    for l in txt.split('\n'):
        if '.section' in l and '.text.' in l:
            inside = kern in l
        m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1), int(m.group(2)))
            continue
        m = re.search(r'/\*([0-9a-f]{4,})\*/\s+(\S.*?);', l)
        if m and inside:
            line_of[int(m.group(1), 16)] = cur

out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True,
                     text=True).stdout.split('\n')
rows = list(csv.reader(out[1:]))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[1:] if len(r) == len(hdr)]
base = int(data[0][0], 16)
cols = ['Instructions Executed', '# Samples', 'L2 Theoretical Sectors Global', 'stall_long_sb', 'stall_short_sb', 'stall_wait',
        'stall_mio', 'stall_lg', 'stall_math', 'stall_barrier']
agg, tot = {}, [0.0] * len(cols)
for r in data:
    key = line_of.get(int(r[0], 16) - base, ('?', -1))
    a = agg.setdefault(key, [0.0] * len(cols))
    for j, c in enumerate(cols):
        v = float(r[ix[c]] or 0) if c in ix else 0.0
        a[j] += v
        tot[j] += v
src = {}
for f in os.listdir(CSRC):
    src[f] = open(os.path.join(CSRC, f), errors='replace').read().split('\n')
print('%s: %d SASS instructions, %d mapped to lines; %.4g warp instructions, %d samples' %
      (kern, len(data), sum(1 for r in data if (int(r[0], 16) - base) in line_of), tot[0], tot[1]))
print('  inst%  smpl%  sect%   long short  wait   mio    lg  math   bar   line')
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    f, ln = key
    text = src.get(f, [''] * 99999)[ln - 1].strip()[:80] if ln > 0 and f in src and ln <= len(src[f]) else ''
    pc = lambda j: 100.0 * a[j] / tot[j] if tot[j] else 0.0
    print('%6.1f %6.1f %6.1f  ' % (pc(0), pc(1), pc(2)) + ' '.join('%5.1f' % (100.0 * a[j] / max(tot[1], 1)) for j in range(3, len(cols))) +
          '   %s:%d  %s' % (f, ln, text))
