"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total and share."""
import csv, re, sys, collections
rows = [r for r in csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"'))]
tot = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
    tot[name][0] += 1
    tot[name][1] += float(r["Metric Value"].replace(",", "")) / 1e3
allus = sum(v[1] for v in tot.values())
print(f"{'kernel':48s} {'launches':>8s} {'total_us':>10s} {'avg_us':>8s} {'share':>6s}")
for k, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:48s} {n:8d} {us:10.1f} {us / n:8.2f} {us / allus:6.1%}")
print(f"{'TOTAL':48s} {sum(v[0] for v in tot.values()):8d} {allus:10.1f}")
