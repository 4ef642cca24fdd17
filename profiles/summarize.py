#!/usr/bin/env python
"""Condense gpurun_out/*.ncu-rep + launches_*.csv into the committed summaries in profiles/.

    python profiles/summarize.py r1

Reads (here, no GPU needed):  ncu -i <rep> --page raw --csv
Writes: profiles/ncu_<tag>_summary.md, profiles/launches_<tag>.csv (copy), profiles/traffic.json
"""
import collections
import csv
import glob
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_barrier.ratio",
    "smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_lg_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_mio_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_not_selected.ratio",
    "smsp__average_warp_latency_issue_stalled_wait.ratio",
]
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def raw(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    head, units = rows[0], rows[1]
    return [dict(zip(head, r)) for r in rows[2:]], dict(zip(head, units))


def to_bytes(val, unit):
    return float(val.replace(",", "")) * UNIT.get(unit, 1)


def main(tag):
    lines = [f"# ncu summaries, tag `{tag}`", "",
             "Produced by `profiles/summarize.py` from `ncu --set full --clock-control none` captures "
             "(`profiles/capture.sh`).  Durations under ncu are cold-cache and serialised; bench numbers "
             "come from `bench.py` only.", ""]
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
    traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
    for rep in sorted(glob.glob(os.path.join(OUT, f"prof_{tag}_*.ncu-rep"))):
        name = os.path.basename(rep)[len(f"prof_{tag}_"):-len(".ncu-rep")]
        kernels, units = raw(rep)
        for k in kernels:
            lines += [f"## {name}: `{k.get('Kernel Name', '?')}`", "", "| metric | value | unit |", "|---|---|---|"]
            for key in KEYS:
                if key in k:
                    lines.append(f"| {key} | {k[key]} | {units.get(key, '')} |")
            rd = to_bytes(k["dram__bytes_read.sum"], units["dram__bytes_read.sum"])
            wr = to_bytes(k["dram__bytes_write.sum"], units["dram__bytes_write.sum"])
            lines += [f"| **dram traffic per launch** | {rd + wr:.0f} | byte |", ""]
            traffic[f"{name}_dram_bytes_per_launch"] = rd + wr
            traffic[f"{name}_tag"] = tag
    lst = os.path.join(OUT, f"launches_{tag}.csv")
    if os.path.exists(lst):
        shutil.copy(lst, os.path.join(ROOT, "profiles", f"launches_{tag}.csv"))
        rows = list(csv.reader(open(lst)))
        hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
        kn, mv = rows[hi].index("Kernel Name"), rows[hi].index("Metric Value")
        agg = collections.defaultdict(lambda: [0, 0.0])
        for r in rows[hi + 1:]:
            try:
                agg[r[kn]][0] += 1
                agg[r[kn]][1] += float(r[mv].replace(",", ""))
            except (ValueError, IndexError):
                pass
        tot = sum(v[1] for v in agg.values())
        lines += [f"## launch list (`launches_{tag}.csv`, `bench.py --steps 2 --warmup 3 --legs n,r`)", "",
                  "| kernel | launches | total ms | avg us | share |", "|---|---|---|---|---|"]
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            lines.append(f"| `{k[:90]}` | {v[0]} | {v[1] / 1e6:.3f} | {v[1] / v[0] / 1e3:.1f} | {100 * v[1] / tot:.1f}% |")
        lines.append("")
    bench = os.path.join(OUT, f"bench_{tag}.json")
    if os.path.exists(bench):
        shutil.copy(bench, os.path.join(ROOT, "profiles", f"bench_{tag}.json"))
    with open(os.path.join(ROOT, "profiles", f"ncu_{tag}_summary.md"), "w") as fh:
        fh.write("\n".join(lines))
    with open(traffic_path, "w") as fh:
        json.dump(traffic, fh, indent=1, sort_keys=True)
    print("\n".join(lines[-25:]))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "r1")
