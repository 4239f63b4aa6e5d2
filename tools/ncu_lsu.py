"""Per-opcode shared-memory wavefronts / global tag requests of an `ncu --page source --csv` dump, per unit of work
(usage: ncu_lsu.py dump.csv [units]; units = e.g. warps x groups, so that the numbers read as "per warp per group")."""
import csv
import re
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {k: i for i, k in enumerate(rows[hi])}
data = rows[hi + 1:]


def num(r, k):
    try:
        return int(r[col[k]] or 0)
    except (ValueError, IndexError):
        return 0


keys = ("Instructions Executed", "L1 Wavefronts Shared", "L1 Wavefronts Shared Excessive", "L1 Tag Requests Global", "# Samples")
g = defaultdict(lambda: [0] * len(keys))
for r in data:
    if len(r) <= col["Source"]:
        continue
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[col["Source"]].strip())
    op = m.group(2) if m else "?"
    base = op.split(".")[0]
    key = op if base in ("LDS", "STS", "LDG", "STG", "LDL", "STL", "SHFL", "ATOMS", "RED") else base
    for i, k in enumerate(keys):
        g[key][i] += num(r, k)
tot = [sum(e[i] for e in g.values()) for i in range(len(keys))]
print(f"{'total':28s} exec {tot[0] / units:9.1f} shared_wf {tot[1] / units:8.1f} excess {tot[2] / units:7.1f} global_tags {tot[3] / units:7.1f} samples {tot[4]}")
for k, e in sorted(g.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[3]) if len(sys.argv) > 3 else 16]:
    print(f"{k:28s} exec {e[0] / units:9.1f} shared_wf {e[1] / units:8.1f} excess {e[2] / units:7.1f} global_tags {e[3] / units:7.1f} samples {e[4]}")
