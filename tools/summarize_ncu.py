#!/usr/bin/env python
"""Selected raw-page metrics of an `ncu --set full` report as a small tracked text file.
usage: python tools/summarize_ncu.py gpurun_out/r02_field_tile.ncu-rep profiles/r02_field_tile_ncu_full.txt "<command line>" """
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(
    r"^(gpu__time_duration.sum|dram__bytes_(read|write).sum|gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed|"
    r"dram__throughput.avg.pct_of_peak_sustained_elapsed|"
    r"launch__(grid_size|block_size|registers_per_thread|waves_per_multiprocessor|occupancy_limit_\w+|cluster_size|shared_mem_per_block_dynamic)|"
    r"lts__t_sector_hit_rate.pct|lts__t_bytes.sum|sm__cycles_(active.avg|active.min|active.max|elapsed.avg.per_second)|"
    r"sm__inst_executed.avg.per_cycle_active|sm__inst_executed_pipe_(alu|fma|fmaheavy|fmalite|xu|lsu|uniform)\.avg\.pct_of_peak_sustained_active|"
    r"sm__pipe_(fma|fmaheavy|alu|xu)_cycles_active.avg.pct_of_peak_sustained_(active|elapsed)|"
    r"sm__throughput.avg.pct_of_peak_sustained_elapsed|sm__warps_active.avg.per_cycle_active|"
    r"smsp__average_warps_issue_stalled_\w+_per_issue_active.ratio|smsp__inst_executed.sum|smsp__issue_active.avg.pct_of_peak_sustained_active|"
    r"l1tex__data_pipe_lsu_wavefronts_mem_shared.sum|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum)$")

rep, out = sys.argv[1], sys.argv[2]
cmd = sys.argv[3] if len(sys.argv) > 3 else ""
raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[-1]
name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else ""
with open(out, "w") as f:
    f.write(f"# {cmd}\n# kernel: {name}\n# metric, unit, value (ncu --page raw; one launch, cold caches, serialised)\n")
    for h, u, v in zip(hdr, units, vals):
        if KEEP.match(h):
            f.write(f"{h},{u},{v}\n")
print("wrote", out, name[:80])
