"""DRAM traffic of the update kernel per launch, from an ncu metrics-only pass over EVERY launch of the kernel in one bench
command:
    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_gemm_tma \
        --csv --log-file <csv> python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-also
Writes profiles/ncu_gemm_traffic.json (read by bench.py -> roofline.traffic) and a text summary.
Usage: python tools/ncu_traffic.py <csv> <workload> <summary.txt> <launches per factorization> <operand bytes per factorization>"""
import csv
import io
import json
import sys

src, workload, out_txt, per_step, operand = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), float(sys.argv[5])
lines = [l for l in open(src) if not l.startswith("==")]
rd = wr = ms = 0.0
n = 0
mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
tmul = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}
for row in csv.DictReader(io.StringIO("".join(lines))):
    name, unit = row.get("Metric Name"), row.get("Metric Unit")
    v = float(row["Metric Value"].replace(",", ""))
    if name == "dram__bytes_read.sum":
        rd += v * mult[unit]; n += 1
    elif name == "dram__bytes_write.sum":
        wr += v * mult[unit]
    elif name == "gpu__time_duration.sum":
        ms += v * tmul[unit]
steps = n / per_step
per_launch = (rd + wr) / max(n, 1)
txt = [f"# {src}: ncu metrics-only pass over every launch of the update kernel ({workload})",
       f"launches captured: {n} = {steps:.2f} factorizations x {per_step} launches",
       f"DRAM read {rd / 1e9:.2f} GB + write {wr / 1e9:.2f} GB = {(rd + wr) / 1e9:.2f} GB -> {(rd + wr) / max(steps, 1e-9) / 1e9:.2f} GB per factorization, {per_launch / 1e6:.2f} MB per launch",
       f"operand bytes if every operand moved exactly once: {operand / 1e9:.2f} GB per factorization ({operand / per_step / 1e6:.2f} MB per launch)",
       f"traffic / that floor: {(rd + wr) / max(steps, 1e-9) / operand:.2f}x",
       f"kernel time under ncu (cold cache, serialised): {ms / max(steps, 1e-9):.1f} ms per factorization"]
open(out_txt, "w").write("\n".join(txt) + "\n")
json.dump({"workload": workload, "kernel": "k_gemm_tma", "launches_captured": n, "dram_bytes_per_launch": per_launch,
           "operand_bytes_per_launch_if_moved_once": operand / per_step, "ratio": (rd + wr) / max(steps, 1e-9) / operand,
           "source": f"{out_txt} (ncu metrics-only pass over all {n} launches of the bench command, mean per launch)"},
          open("profiles/ncu_gemm_traffic.json", "w"), indent=1)
print("\n".join(txt))
