"""Hot SASS instructions of an `ncu --page source --csv` dump: samples, dominant stall reason, and a running index so
that regions (producer / MMA / epilogue phases) can be told apart.  usage: ncu_sass_hot.py dump.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hdr_i]
col = {k: i for i, k in enumerate(h)}
stall_cols = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
data = []
for idx, r in enumerate(rows[hdr_i + 1:]):
    if len(r) < len(h):
        continue
    try:
        s = int(r[col["# Samples"]])
    except ValueError:
        continue
    st = sorted(((int(r[col[k]] or 0), k) for k in stall_cols), reverse=True)[:2]
    data.append((s, idx, r[col["Source"]].strip(), st, int(r[col["Instructions Executed"]] or 0)))
tot = sum(d[0] for d in data)
print("total samples", tot, "instructions", len(data))
# region summary: split at big gaps is manual; print cumulative samples per 100 instructions
for a in range(0, len(data), 100):
    blk = data[a:a + 100]
    print(f"[{a:5d}-{a + len(blk) - 1:5d}] samples {sum(d[0] for d in blk):7d}  exec {sum(d[4] for d in blk):10d}")
print()
for s, idx, src, st, ex in sorted(data, reverse=True)[:top]:
    print(f"{s:7d} {100.0 * s / tot:5.1f}%  #{idx:5d} exec {ex:9d}  {src[:70]:70s} {st[0][1]}={st[0][0]} {st[1][1]}={st[1][0]}")
