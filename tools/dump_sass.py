#!/usr/bin/env python
"""Dump per-kernel SASS listings + a mnemonic histogram (the evidence for tcgen05 / TMA use) into profiles/sass/."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "myfyp_b200", "build", "obj")
OUT = os.path.join(ROOT, "profiles", "sass")
EVIDENCE = re.compile(r"\b(UTC[A-Z0-9]*MMA[A-Z0-9_.]*|LDTM[A-Z0-9_.]*|STTM[A-Z0-9_.]*|UTMALDG[A-Z0-9_.]*|UTMASTG[A-Z0-9_.]*|UBLKCP[A-Z0-9_.]*|UTCBAR[A-Z0-9_.]*|UTCATOMSWS[A-Z0-9_.]*|SYNCS[A-Z0-9_.]*|HMMA[A-Z0-9_.]*)")


def main():
    os.makedirs(OUT, exist_ok=True)
    summary = []
    for obj in sorted(os.listdir(OBJ)):
        if not obj.endswith(".o"):
            continue
        sass = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, obj)], capture_output=True, text=True).stdout
        cur, body = None, {}
        for line in sass.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                cur = m.group(1)
                body[cur] = []
            elif cur:
                body[cur].append(line)
        for fn, lines in body.items():
            demangled = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip().split("(")[0].replace("void ", "").replace("b2::", "")
            short = re.sub(r"[^A-Za-z0-9_]+", "_", demangled).strip("_")[:60]
            ops = collections.Counter()
            for ln in lines:
                m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", ln)
                if m:
                    ops[m.group(1).split(".")[0]] += 1
            ev = collections.Counter(x for ln in lines for x in EVIDENCE.findall(ln))
            keep_full = bool(ev) or len(lines) < 400
            with open(os.path.join(OUT, short + ".sass"), "w") as f:
                f.write(f"// {demangled}   ({obj})  {sum(ops.values())} instructions\n// opcode histogram: {dict(ops.most_common(14))}\n")
                if ev:
                    f.write(f"// Blackwell evidence: {dict(ev)}\n")
                f.write("\n".join(lines if keep_full else lines[:200] + ["// ... truncated (pure ALU/LSU kernel) ..."]))
            summary.append((demangled, obj, sum(ops.values()), dict(ev)))
    with open(os.path.join(OUT, "README.md"), "w") as f:
        f.write("# SASS listings (cuobjdump -sass, sm_100a)\n\n| kernel | object | instructions | tcgen05 / TMA / mbarrier mnemonics |\n|---|---|---|---|\n")
        for d, o, n, ev in summary:
            f.write(f"| `{d}` | {o} | {n} | {ev if ev else '-'} |\n")
    print(f"{len(summary)} kernels")


if __name__ == "__main__":
    main()
