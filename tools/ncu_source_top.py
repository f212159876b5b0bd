"""Top stalled source lines / SASS instructions of one kernel from `ncu -i rep --page source --csv [--launch-skip N --launch-count 1]`."""
import csv, sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if "# Samples" in r)
hdr = rows[hi]
si, src = hdr.index("# Samples"), hdr.index("Source")
def num(x):
    try: return int(float(x))
    except ValueError: return 0
data = [r for r in rows[hi + 1:] if len(r) > si and "# Samples" not in r]
tot = sum(num(r[si]) for r in data)
print("kernel:", rows[0][1][:120] if rows[0] else "?", "| total samples", tot)
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
for r in sorted(data, key=lambda r: -num(r[si]))[:n]:
    st = sorted(((hdr[i], num(r[i])) for i in stall if num(r[i]) > 0), key=lambda kv: -kv[1])[:3]
    print(f"{num(r[si]):6d} {100 * num(r[si]) / max(tot, 1):5.1f}%  {r[src][:100]:100s} {st}")
