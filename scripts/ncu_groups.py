"""Print per-kernel durations from an `ncu --metrics gpu__time_duration.sum --csv` log, one line per group of launches
between L2-flush fills (scripts/bench_search.py --profile)."""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, mi = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Name')
out = []
for r in rows[1:]:
    if r[mi] != 'gpu__time_duration.sum':
        continue
    name, v = r[ki], float(r[vi].replace(',', ''))
    if 'FillFunctor' in name or 'elementwise' in name:
        if out:
            print(' | '.join(out))
            out = []
        continue
    out.append(f"{name.split('::')[-1].split('(')[0][:28]}={v / 1000:.1f}")
if out:
    print(' | '.join(out))
