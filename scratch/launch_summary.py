"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel count, avg us, share."""
import csv, sys, collections, re
path = sys.argv[1]
rows = []
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
tot = collections.defaultdict(lambda: [0, 0.0])
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    if unit in ("nsecond", "ns"): v /= 1e3
    elif unit in ("msecond", "ms"): v *= 1e3
    elif unit in ("second", "s"): v *= 1e6
    tot[name][0] += 1; tot[name][1] += v
s = sum(v[1] for v in tot.values())
for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:70]:70s} n={n:4d} avg={t/n:10.1f} us total={t:10.1f} us share={100*t/s:5.1f}%")
