#!/usr/bin/env python
"""Summarise an `ncu --set full` report into one line per launch (used to fill profiles/).

    ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/raw.csv ; python tools/ncu_summary.py /tmp/raw.csv
"""
import csv
import sys

COLS = [
    ("time_us", "gpu__time_duration.sum", 1e-3),
    ("dram_rd_MB", "dram__bytes_read.sum", None),
    ("dram_wr_MB", "dram__bytes_write.sum", None),
    ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("tensor_pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 1),
    ("sm_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("warps_act_pct", "sm__warps_active.avg.pct_of_peak_sustained_active", 1),
    ("regs", "launch__registers_per_thread", 1),
    ("smem_dyn_KB", "launch__shared_mem_per_block_dynamic", None),
    ("l1_hit_pct", "l1tex__t_sector_hit_rate.pct", 1),
    ("l2_hit_pct", "lts__t_sector_hit_rate.pct", 1),
]
UNIT = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
UNIT_KB = {"byte": 1 / 1024, "Kbyte": 1.0, "Mbyte": 1024.0}
TIME = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in rows[2:]:
        d = {"id": r[col["ID"]], "kernel": r[col["Kernel Name"]].replace("<unnamed>::", "")[:60],
             "grid": r[col["Grid Size"]], "block": r[col["Block Size"]]}
        for name, metric, scale in COLS:
            if metric not in col:
                d[name] = ""
                continue
            try:
                v = float(r[col[metric]].replace(",", ""))
            except ValueError:
                d[name] = ""
                continue
            u = units[col[metric]]
            if name == "time_us":
                v *= TIME.get(u, 1.0)
            elif name.endswith("_MB"):
                v *= UNIT.get(u, 1.0)
            elif name.endswith("_KB"):
                v *= UNIT_KB.get(u, 1.0)
            d[name] = round(v, 2)
        out.append(d)
    keys = ["id", "kernel", "grid", "block"] + [c[0] for c in COLS]
    w = csv.DictWriter(sys.stdout, fieldnames=keys)
    w.writeheader()
    for d in out:
        w.writerow(d)


if __name__ == "__main__":
    main(sys.argv[1])
