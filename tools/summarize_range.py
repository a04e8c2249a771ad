"""Turn the ncu range-replay csv files of tools/dram_range.py into profiles/latest_ncu.json keys.
usage: python tools/summarize_range.py <tag> <launches> observe=<csv> step=<csv>"""
import csv
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def read(path):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith('==')) if r]
    hdr = rows[0]
    im, iv, iu = hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
    out = {}
    for r in rows[1:]:
        if len(r) <= iv:
            continue
        v = float(r[iv].replace(',', ''))
        u = r[iu]
        scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'ns': 1e-9, 'us': 1e-6, 'ms': 1e-3, 's': 1}.get(u, 1)
        out[r[im]] = out.get(r[im], 0.0) + v * scale
    return out


def main():
    tag, launches = sys.argv[1], int(sys.argv[2])
    p = os.path.join(REPO, 'profiles', 'latest_ncu.json')
    d = json.load(open(p))
    notes = []
    for spec in sys.argv[3:]:
        k, path = spec.split('=')
        if not os.path.exists(path):
            continue
        m = read(path)
        b = (m.get('dram__bytes_read.sum', 0) + m.get('dram__bytes_write.sum', 0)) / launches
        d['k_%s_dram_bytes_range_per_launch' % k] = b
        d['k_%s_range_seconds_per_launch_under_ncu' % k] = m.get('gpu__time_duration.sum', 0) / launches
        notes.append('k_%s: %.1f MB read + %.1f MB written per launch' % (
            k, m.get('dram__bytes_read.sum', 0) / launches / 1e6, m.get('dram__bytes_write.sum', 0) / launches / 1e6))
    key = 'range_source_two_groups' if any('two_groups' in s_ for s_ in sys.argv[3:]) else 'range_source'
    d[key] = ('profiles/%s ncu --replay-mode range --cache-control none over %d back-to-back launches '
              '(tools/dram_range.py): ' % (tag, launches)) + '; '.join(notes)
    json.dump(d, open(p, 'w'), indent=1)
    print(d[key])
    return
    json.dump(d, open(p, 'w'), indent=1)
    print(d['range_source'])


if __name__ == '__main__':
    main()
