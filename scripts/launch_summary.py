"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python scripts/launch_summary.py launches.csv"""
import collections, csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row["Metric Unit"], 1.0)
    a = agg[row["Kernel Name"][:70]]
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("%8s %12s %10s %6s  kernel" % ("launches", "total_ms", "avg_us", "share"))
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%8d %12.3f %10.1f %5.1f%%  %s" % (n, t / 1e3, t / n, 100 * t / tot, k))
print("%8d %12.3f" % (sum(a[0] for a in agg.values()), tot / 1e3))
