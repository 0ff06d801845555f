#!/usr/bin/env python
"""SASS census of libeeg_b200.so: per kernel, how many tensor-core / TMA / packed-FMA instructions the shipped binary holds
(cuobjdump -sass; no GPU needed).  Writes profiles/<round>_sass_census.txt.

    python tools/sass_census.py [out.txt]
UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st (TMEM), UBLKCP = cp.async.bulk (TMA engine, non-tensor form),
UTMALDG/UTMASTG = cp.async.bulk.tensor (tensor-map TMA), HMMA = mma.sync, FFMA2 = packed fp32x2 FMA."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "energy_equiv_lattice_gnn_b200", "libeeg_b200.so")
OPS = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "HMMA", "FFMA2", "FMUL2", "FFMA", "LDGSTS", "LDS", "STS", "ATOMG", "RED", "SHFL"]


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_census.txt")
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip() or n  # noqa: E731
    cur, counts = None, collections.OrderedDict()
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            op = m.group(1)
            counts[cur][op] += 1
            counts[cur]["_total"] += 1
    rows = []
    for fn, c in counts.items():
        name = demangle(fn)
        name = re.sub(r"\(anonymous namespace\)::", "", name)
        name = re.sub(r"\(.*", "", name)[:60]
        rows.append((name, c))
    rows.sort(key=lambda r: (-r[1]["UTCHMMA"], -r[1]["HMMA"], -r[1]["_total"]))
    with open(out, "w") as f:
        f.write("SASS census of energy_equiv_lattice_gnn_b200/libeeg_b200.so (cuobjdump -sass, sm_100a)\n")
        f.write(f"{'kernel':60s} {'instr':>7s} " + " ".join(f"{o:>7s}" for o in OPS) + "\n")
        tot = collections.Counter()
        for name, c in rows:
            f.write(f"{name:60s} {c['_total']:7d} " + " ".join(f"{c[o]:7d}" for o in OPS) + "\n")
            tot.update(c)
        f.write(f"{'TOTAL':60s} {tot['_total']:7d} " + " ".join(f"{tot[o]:7d}" for o in OPS) + "\n")
    print(open(out).read()[:3000])


if __name__ == "__main__":
    main()
