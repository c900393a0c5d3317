"""Summarise an `ncu --csv --metrics gpu__time_duration.sum` launch list by kernel name."""
import csv, sys, collections, re
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
r = csv.DictReader(lines)
agg = collections.OrderedDict()
tot = 0.0
for row in r:
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", row["Kernel Name"])
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += ns; tot += ns
print(f"total {tot/1e6:.3f} ms over {sum(a[0] for a in agg.values())} launches")
print(f"{'kernel':70s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
for name, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name[:70]:70s} {n:8d} {ns/1e6:10.3f} {100*ns/tot:6.1f}%")
