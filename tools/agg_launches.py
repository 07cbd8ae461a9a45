"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    ms = v / 1e6 if u.startswith("ns") else (v / 1e3 if u.startswith("us") else v)
    name = re.sub(r"\(.*", "", row["Kernel Name"])[:80]
    agg[name][0] += 1
    agg[name][1] += ms
tot = sum(v[1] for v in agg.values())
nsteps = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
print("%d launches, %.2f ms total, %.2f ms per step (%g steps captured)" % (sum(v[0] for v in agg.values()), tot, tot / nsteps, nsteps))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 25]:
    print("%8.2f ms/step %5.1f%% %6.1f launches/step  %s" % (v[1] / nsteps, 100 * v[1] / tot, v[0] / nsteps, k))
