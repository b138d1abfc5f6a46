#!/bin/bash
# 1-GPU visit: parity tests, then short benches
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
for wl in "$@"; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err
  echo "bench $wl rc=$?"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_$wl.json"))
    print(d["config"]["workload"], "ms", round(d["ms_per_step"],2), "GF/s", round(d["value"]), "gemm frac", d["roofline"]["frac"], d["kernel_ms"])
except Exception as e: print("no json", e)
PY
  tail -3 gpurun_out/bench_$wl.err
done
