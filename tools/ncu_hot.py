#!/usr/bin/env python
"""Top stall sites of one kernel in an ncu report (SASS view of `--page source`).

    python tools/ncu_hot.py REPORT.ncu-rep KERNEL_REGEX [launch_skip] [top_n]
"""
import csv
import io
import subprocess
import sys


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    skip = sys.argv[3] if len(sys.argv) > 3 else "0"
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    view = sys.argv[5] if len(sys.argv) > 5 else "sass"
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", view, "-k", f"regex:{rx}",
                          "--launch-count", "1"] + (["--launch-skip", skip] if skip != "0" else []), capture_output=True, text=True)
    if not out.stdout.strip():
        sys.exit(out.stderr)
    out = out.stdout
    lines = out.splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"Address"') or l.startswith('"#"') or l.startswith('"Line'))
    print(lines[0][:200])
    rows = list(csv.reader(io.StringIO("\n".join(lines[start:]))))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    samp = col.get("# Samples", col.get("Warp Stall Sampling (All Samples)"))
    stall_cols = [(h, i) for h, i in col.items() if h.startswith("stall_") and "Not Issued" not in h]
    body = [r for r in rows[1:] if len(r) > samp and (r[samp] or "0").isdigit()]
    total = sum(int(r[samp] or 0) for r in body)
    print("total samples", total, "instructions", len(body))
    agg = {}
    for r in body:
        for h, i in stall_cols:
            agg[h] = agg.get(h, 0) + int(r[i] or 0)
    print("stall totals:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
    idx = sorted(range(len(body)), key=lambda i: -int(body[i][samp] or 0))[:top]
    for i in sorted(idx):
        r = body[i]
        st = {h[6:]: int(r[c] or 0) for h, c in stall_cols if int(r[c] or 0)}
        ex = r[col["Instructions Executed"]] if "Instructions Executed" in col else ""
        print(f"{i:5d} {int(r[samp] or 0):6d} {100.0 * int(r[samp] or 0) / max(total, 1):5.1f}%  ex={ex:>9s} {r[1][:90]:90s} {st}")


if __name__ == "__main__":
    main()
