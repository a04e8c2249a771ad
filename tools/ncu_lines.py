"""Per-source-line instruction counts / stall samples from `ncu --page source --csv --print-source cuda,sass`."""
import csv
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
fname = None
hdr = None
out = []
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        fname = r[1].split('/')[-1]
        continue
    if r[0] == 'Line No':
        hdr = r
        continue
    if hdr and r[0] not in ('', 'Function Name') and r[0].isdigit():
        d = dict(zip(hdr[4:], r[4:]))
        try:
            out.append((int(d['Instructions Executed']), int(d['# Samples']), int(d['Thread Instructions Executed']),
                        fname, int(r[0]), r[1].strip()[:110]))
        except (ValueError, KeyError):
            pass
tot = sum(o[0] for o in out)
tots = sum(o[1] for o in out)
print('total warp-inst %d  samples %d' % (tot, tots))
out.sort(reverse=True)
for o in out[:top]:
    print('%5.1f%% inst %5.1f%% smp thr/inst %4.1f  %s:%d  %s' % (100.0 * o[0] / tot, 100.0 * o[1] / max(tots, 1),
                                                                o[2] / max(o[0], 1), o[3], o[4], o[5]))

# per-file and per-phase (line range of dca_kernels.cuh) shares
from collections import defaultdict
byfile = defaultdict(int)
for o in out:
    byfile[o[3]] += o[0]
print('--- by file')
for k, v in sorted(byfile.items(), key=lambda kv: -kv[1]):
    print('%5.1f%%  %s' % (100.0 * v / tot, k))
if len(sys.argv) > 3:
    print('--- by range of %s' % sys.argv[3])
    ranges = [tuple(map(int, a.split('-'))) for a in sys.argv[4:]]
    for lo, hi in ranges:
        v = sum(o[0] for o in out if o[3] == sys.argv[3] and lo <= o[4] <= hi)
        sm = sum(o[1] for o in out if o[3] == sys.argv[3] and lo <= o[4] <= hi)
        print('%5.1f%% inst %5.1f%% smp  lines %d-%d' % (100.0 * v / tot, 100.0 * sm / tots, lo, hi))
