#!/usr/bin/env python
"""ncu -i <rep> --page raw --csv for several reports -> ONE table, a row per captured kernel, of the metrics that place a
secondary kernel against its bound.  usage: ncu_multi_summary.py "header comment" rep1.ncu-rep rep2.ncu-rep ... > out.csv"""
import csv
import io
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "time"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__registers_per_thread", "regs"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_pct"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_pct"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu_pct"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu_pct"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pct"),
        ("sass__inst_executed_local_loads", "local_ld")]
print("# " + sys.argv[1])
print("kernel," + ",".join(n for _, n in COLS))
for rep in sys.argv[2:]:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        continue
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
        name = d.get("Kernel Name", ("?", ""))[0].split("(")[0]
        out = []
        for key, _ in COLS:
            v, u = d.get(key, ("", ""))
            out.append(f"{v} {u}".strip() if u not in ("", "%", "inst", "register/thread", "block") else v)
        print('"' + name + '",' + ",".join(out))
