"""ncu launch-list CSV (--metrics gpu__time_duration.sum --csv) -> per-kernel table (markdown).
usage: python scripts/launch_summary.py launches.csv [skip_first_n]"""
import collections
import csv
import sys

rows, hdr = [], None
for r in csv.reader(open(sys.argv[1], errors="ignore")):
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr is None or len(r) != len(hdr):
        continue
    d = dict(zip(hdr, r))
    if d.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(d["Metric Value"].replace(",", ""))
    u = d.get("Metric Unit", "ns")
    v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
    rows.append((d["Kernel Name"], v))
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = rows[skip:]
agg = collections.defaultdict(lambda: [0, 0.0])
for k, v in rows:
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v for _, v in rows)
print(f"{len(rows)} launches, {tot / 1e3:.2f} ms\n\nkernel | launches | mean us | total ms | share\n---|---|---|---|---")
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:40]:
    print(f"{k[-90:]} | {n} | {t / n:.1f} | {t / 1e3:.2f} | {100 * t / tot:.1f}%")
