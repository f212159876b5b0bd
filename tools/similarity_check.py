"""Same-named files of this package vs the reference: difflib character ratio (a self-check that nothing drifted towards a copy)."""
import difflib, os, sys
ref_root = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/p2pfl"
my_root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "myfyp_b200")
ref = {}
for d, _, fs in os.walk(ref_root):
    for f in fs:
        if f.endswith(".py"):
            ref.setdefault(f, []).append(os.path.join(d, f))
out = []
for d, _, fs in os.walk(my_root):
    for f in fs:
        if f.endswith(".py") and f in ref:
            a = open(os.path.join(d, f)).read()
            if len(a) < 50:
                continue
            for rp in ref[f]:
                b = open(rp).read()
                out.append((difflib.SequenceMatcher(None, a, b, autojunk=False).ratio(), os.path.join(d, f)[len(my_root) + 1:], rp[len(ref_root) + 1:]))
for r, a, b in sorted(out, reverse=True)[: int(sys.argv[2]) if len(sys.argv) > 2 else 15]:
    print(f"{r:.2f} {a}  vs  {b}")
