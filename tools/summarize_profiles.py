#!/usr/bin/env python
"""Turn the raw ncu outputs of tools/profile_round.sh (gpurun_out/) into the tracked summaries under profiles/:
  <tag>_launches_bench.csv      per-kernel aggregate of the launch list + the first launches verbatim
  <tag>_cpairs_kernel_ncu_full.txt   selected raw-page metrics of the one `--set full` capture
  traffic.json                  DRAM bytes per launch of the dominant kernel (read by bench.py)
usage: python tools/summarize_profiles.py r01b"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"

KEEP = re.compile(
    r"^(gpu__time_duration.sum|dram__bytes_(read|write).sum|gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed|"
    r"launch__(grid_size|block_size|registers_per_thread|waves_per_multiprocessor|occupancy_limit_\w+)|"
    r"lts__t_sector_hit_rate.pct|idc__request_hit_rate.pct|sm__cycles_(active.avg|elapsed.avg.per_second)|"
    r"sm__inst_executed.avg.per_cycle_active|sm__inst_executed_pipe_(alu|fma|fmaheavy|fmalite|xu|lsu|uniform)\.avg\.pct_of_peak_sustained_active|"
    r"sm__pipe_(fma|fmaheavy|alu|xu)_cycles_active.avg.pct_of_peak_sustained_(active|elapsed)|"
    r"sm__throughput.avg.pct_of_peak_sustained_elapsed|sm__warps_active.avg.per_cycle_active|"
    r"smsp__average_warps_issue_stalled_\w+_per_issue_active.ratio|smsp__inst_executed.sum|"
    r"l1tex__data_pipe_lsu_wavefronts_mem_shared.sum|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum)$")


def launches():
    rows = [r for r in csv.reader(open(os.path.join(OUT, "launches.csv"))) if len(r) > 10]
    hdr, body = rows[0], rows[1:]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    gi, bi = hdr.index("Grid Size"), hdr.index("Block Size")
    agg = collections.OrderedDict()
    for r in body:
        name = re.sub(r"\(.*", "", r[ki])
        us = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
        a = agg.setdefault(name, {"n": 0, "total_us": 0.0, "min_us": 1e30, "max_us": 0.0, "grid": r[gi], "block": r[bi]})
        a["n"] += 1
        a["total_us"] += us
        a["min_us"] = min(a["min_us"], us)
        a["max_us"] = max(a["max_us"], us)
    total = sum(a["total_us"] for a in agg.values())
    path = os.path.join(PROF, f"{tag}_launches_bench.csv")
    with open(path, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv python bench.py --steps 1 --warmup 3 "
                "--no-cpu-baseline --no-extras\n")
        f.write(f"# {len(body)} launches; per-kernel aggregate first, then the first 40 launches verbatim\n")
        f.write("kernel,launches,total_us,share,mean_us,min_us,max_us,grid,block\n")
        for k, a in agg.items():
            f.write(f"\"{k}\",{a['n']},{a['total_us']:.1f},{a['total_us'] / total:.5f},{a['total_us'] / a['n']:.2f},"
                    f"{a['min_us']:.2f},{a['max_us']:.2f},\"{a['grid']}\",\"{a['block']}\"\n")
        f.write("# ---- first launches\n")
        f.write(",".join(hdr) + "\n")
        for r in body[:40]:
            f.write(",".join('"%s"' % c for c in r) + "\n")
    print("wrote", path)
    return agg


def full():
    rep = os.path.join(OUT, "cpairs_full.ncu-rep")
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units, val = rows[0], rows[1], rows[2]
    path = os.path.join(PROF, f"{tag}_cpairs_kernel_ncu_full.txt")
    got = {}
    with open(path, "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on -k regex:cpairs_kernel -s 1700 -c 1 "
                "python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras   (raw page, selected metrics)\n")
        f.write(f"kernel\t\t{val[hdr.index('Kernel Name')]}\n")
        for h, u, v in sorted(zip(hdr, units, val)):
            if KEEP.match(h):
                f.write(f"{h}\t{u}\t{v}\n")
                got[h] = (u, v)
    print("wrote", path)
    return got


def extra(rep_name, out_name, note):
    """Selected raw-page metrics of one additional `--set full` capture (same metric list)."""
    rep = os.path.join(OUT, rep_name)
    if not os.path.exists(rep):
        print("missing", rep)
        return
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units, val = rows[0], rows[1], rows[2]
    path = os.path.join(PROF, f"{tag}_{out_name}_ncu_full.txt")
    with open(path, "w") as f:
        f.write(f"# {note}   (raw page, selected metrics)\n")
        f.write(f"kernel\t\t{val[hdr.index('Kernel Name')]}\n")
        for h, u, v in sorted(zip(hdr, units, val)):
            if KEEP.match(h) or h.startswith("smsp__inst_executed_op") or h in ("l1tex__t_bytes.sum", "lts__t_bytes.sum"):
                f.write(f"{h}\t{u}\t{v}\n")
    print("wrote", path)


def to_bytes(u, v):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]


if __name__ == "__main__":
    launches()
    g = full()
    tpath = os.path.join(PROF, "traffic.json")
    try:
        t = json.load(open(tpath))
    except (OSError, ValueError):
        t = {}
    bench = json.loads(open(os.path.join(OUT, "bench_short.json")).read().strip().splitlines()[-1])
    pts = bench["config"]["targets_per_gpu"]
    t[f"cpairs_kernel_field_896x{pts}"] = {
        "dram_bytes_read": to_bytes(*g["dram__bytes_read.sum"]), "dram_bytes_write": to_bytes(*g["dram__bytes_write.sum"]),
        "source": f"profiles/{tag}_cpairs_kernel_ncu_full.txt (ncu --set full, one launch = 896 sources x {pts} targets)"}
    json.dump(t, open(tpath, "w"), indent=1)
    print("wrote", tpath)
    extra("flip_full.ncu-rep", "flip_exhaustive_kernel",
          "ncu --set full --clock-control none --import-source on -k regex:flip_exhaustive_kernel -s 18 -c 1 python tools/flip_probe.py  (36 patches, 2^35 states)")
    extra("parse_full.ncu-rep", "parse_lines_kernel",
          "ncu --set full --clock-control none --import-source on -k regex:parse_lines_kernel -s 2 -c 1 python tools/io_gpu_probe.py  (1M points, 124 MB of text)")
    extra("format_full.ncu-rep", "format_values_kernel",
          "ncu --set full --clock-control none --import-source on -k regex:format_values_kernel -s 2 -c 1 python tools/io_gpu_probe.py  (1M points x 6 values)")
    for name in ("bench.json", "flip_probe.json", "io_probe.json"):
        src = os.path.join(OUT, name)
        if os.path.exists(src):
            open(os.path.join(PROF, f"{tag}_{name}"), "w").write(open(src).read())
            print("copied", name)
