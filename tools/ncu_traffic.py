"""Reduce an `ncu --set full` report of k_gemm_dmma launches to profiles/ncu_gemm_traffic.json + a text summary.
Usage: python tools/ncu_traffic.py <report.ncu-rep> <workload> <summary.txt> "<command line that was profiled>" """
import csv
import io
import json
import subprocess
import sys

rep, workload, out_txt, cmdline = sys.argv[1:5]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {k: i for i, k in enumerate(hdr)}


def f(row, key):
    v = row[col[key]].replace(",", "")
    return float(v) if v else 0.0


def to_bytes(v, unit):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__cycles_active.avg",
        "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
        "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"]
dm = [k for k in hdr if "subpipe_dmma_cycles_active" in k]
lines = [f"# ncu --set full --clock-control none --import-source on -k regex:k_gemm_dmma ({cmdline}), 1 B200",
         f"# report: {rep} (scratch); per-launch values of the {len(data)} captured launches"]
tot_bytes, tot_ms = 0.0, 0.0
for k in ["Kernel Name", "Grid Size", "Block Size"] + keys + dm:
    if k not in col:
        continue
    lines.append(f"{k} [{units[col[k]]}]: {[r[col[k]] for r in data]}")
per = []
for r in data:
    b = to_bytes(f(r, "dram__bytes_read.sum"), units[col["dram__bytes_read.sum"]]) + \
        to_bytes(f(r, "dram__bytes_write.sum"), units[col["dram__bytes_write.sum"]])
    per.append(b)
if dm:
    lines.append("DMMA sub-pipe active / SMSP active cycles: " +
                 str([round(f(r, dm[0]) / max(f(r, "smsp__cycles_active.avg"), 1), 3) for r in data]))
lines.append(f"traffic per launch (dram read+write) [MB]: {[round(b / 1e6, 1) for b in per]}")
open(out_txt, "w").write("\n".join(lines) + "\n")
json.dump({"workload": workload, "kernel": "k_gemm_dmma", "launches_captured": len(data),
           "dram_bytes_per_launch": sum(per) / max(len(per), 1),
           "source": f"{out_txt} (ncu --set full, mean of {len(data)} captured launches of the bench command)"},
          open("profiles/ncu_gemm_traffic.json", "w"), indent=1)
print("\n".join(lines)[:3000])
