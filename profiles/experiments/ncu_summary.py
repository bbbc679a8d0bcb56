#!/usr/bin/env python
"""ncu -i <rep> --page raw --csv  ->  'metric,unit,value' rows of the metrics quoted in DESIGN.md (first kernel of the
report).  usage: ncu_summary.py report.ncu-rep "header comment" > profiles/<name>.csv"""
import csv
import io
import re
import subprocess
import sys

rep, note = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
keep = re.compile(r"^(Kernel Name|dram__bytes_(read|write)\.sum$|gpu__time_duration\.sum$|launch__(block_size|grid_size|registers_per_thread|"
                  r"occupancy_limit_\w+|shared_mem_per_block_dynamic|waves_per_multiprocessor)$|sass__inst_executed_local_(loads|stores)$|"
                  r"sm__cycles_elapsed\.avg(\.per_second)?$|sm__inst_executed_pipe_(alu|fma|fmaheavy|fmalite|lsu|xu|uniform)\.avg\.pct_of_peak_sustained_active$|"
                  r"sm__pipe_(fma|fmaheavy|fmalite|alu|xu)_cycles_active\.avg\.pct_of_peak_sustained_active$|sm__throughput\.avg\.pct_of_peak_sustained_elapsed$|"
                  r"sm__warps_active\.avg\.pct_of_peak_sustained_active$|smsp__inst_executed\.sum$|smsp__issue_active\.avg\.pct_of_peak_sustained_active$|"
                  r"smsp__warps_eligible\.avg\.per_cycle_active$|smsp__sass_thread_inst_executed_op_(fadd|ffma|fmul|fp32)_pred_on\.sum(\.per_cycle_elapsed)?$|"
                  r"smsp__sass_thread_inst_executed_ops_fadd_fmul_ffma_pred_on\.sum$|smsp__inst_executed_pipe_(fma|xu|lsu|alu)\.sum$|"
                  r"smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio$|sm__sass_inst_executed_op_shared_(ld|st)\.sum$|"
                  r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum$|sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_active$)")
print("# " + note)
print("metric,unit,value")
for h, u, v in zip(hdr, units, vals):
    if keep.match(h):
        print(f'{h},{u},"{v}"' if "," in v else f"{h},{u},{v}")
