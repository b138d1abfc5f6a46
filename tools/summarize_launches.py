"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (-> profiles/)."""
import collections
import csv
import io
import sys

src, dst = sys.argv[1], sys.argv[2]
lines = [l for l in open(src) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in csv.DictReader(io.StringIO("".join(lines))):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = row["Kernel Name"].split("(")[0].replace("void ", "")
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    v = v / 1e6 if unit == "ns" else v / 1e3 if unit.startswith("us") else v
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
with open(dst, "w") as f:
    f.write(f"# source: {src} (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised: compare SHARES)\n")
    f.write(f"# {' '.join(sys.argv[3:])}\n")
    f.write(f"{'kernel':40s} {'launches':>8s} {'total_ms':>10s} {'avg_us':>9s} {'share':>6s}\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k:40s} {v[0]:8d} {v[1]:10.3f} {v[1] / v[0] * 1e3:9.1f} {v[1] / tot:6.3f}\n")
    f.write(f"{'TOTAL':40s} {sum(v[0] for v in agg.values()):8d} {tot:10.3f}\n")
print(open(dst).read())
