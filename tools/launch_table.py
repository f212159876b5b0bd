"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel counts / time shares, optionally every launch."""
import collections, csv, re, sys

path = sys.argv[1]
every = len(sys.argv) > 2
lines = [l for l in open(path) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
tot, rows = 0.0, []
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", ""))
    v = v / 1000 if row["Metric Unit"] == "ns" else v * 1000 if row["Metric Unit"] == "ms" else v
    name = re.sub(r"^void ", "", re.sub(r"\(.*", "", row["Kernel Name"])).replace("b2::", "")
    agg[name][0] += 1; agg[name][1] += v; tot += v
    rows.append((name, row.get("Grid Size", ""), v))
print(f"{len(rows)} launches, {tot:.1f} us summed\n\n| kernel | launches | us | share |\n|---|---|---|---|")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{k[:80]}` | {c} | {t:.1f} | {100 * t / tot:.1f}% |")
if every:
    print("\n| # | kernel | grid | us |\n|---|---|---|---|")
    for i, (n, g, v) in enumerate(rows):
        print(f"| {i} | `{n[:60]}` | {g} | {v:.1f} |")
