"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and launch count per kernel."""
import csv, sys, collections, re
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]); name = re.sub(r"^void ", "", name)
    if "cub::" in name: name = "cub::" + re.sub(r"<.*", "", name.split("::")[-1])
    t = float(r[14]) * ({"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[13], 1e-6))
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += t
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':50s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:50]:50s} {n:8d} {t:10.3f} {100 * t / tot:6.1f}%")
print(f"{'total':50s} {len(rows):8d} {tot:10.3f}")
