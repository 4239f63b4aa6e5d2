#!/usr/bin/env python
"""Stall samples of one kernel aggregated over instruction-index ranges (SASS view).

    python tools/ncu_regions.py REPORT.ncu-rep KERNEL_REGEX b0,b1,b2,...   # ranges [0,b0) [b0,b1) ...
"""
import csv
import io
import subprocess
import sys


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    bounds = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else []
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", f"regex:{rx}",
                          "--launch-count", "1"], capture_output=True, text=True).stdout
    lines = out.splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"Address"') or l.startswith('"#"'))
    rows = list(csv.reader(io.StringIO("\n".join(lines[start:]))))
    col = {h: i for i, h in enumerate(rows[0])}
    samp = col.get("# Samples", col.get("Warp Stall Sampling (All Samples)"))
    ex = col.get("# Instructions Executed", col.get("Instructions Executed"))
    body = [r for r in rows[1:] if len(r) > samp]
    total = sum(int(r[samp] or 0) for r in body)
    edges = [0] + bounds + [len(body)]
    for a, b in zip(edges[:-1], edges[1:]):
        s = sum(int(r[samp] or 0) for r in body[a:b])
        e = sum(int(r[ex] or 0) for r in body[a:b]) if ex is not None else 0
        print(f"[{a:5d},{b:5d})  samples {s:7d} {100.0 * s / max(total, 1):5.1f}%   warp-instr executed {e}")


if __name__ == "__main__":
    main()
