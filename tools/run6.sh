#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
show() { python -c "
import json,sys
d=json.loads(open('$1').read().strip().splitlines()[-1])
print('$1', round(d['ms_per_step'],2), round(d['value']), d['kernel_ms'], d['roofline']['by_K_class'])"; }
for nb in 256 512 1024; do
  SPCHOL_NB=$nb timeout 900 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_128_nb$nb.json 2> gpurun_out/bench_128_nb$nb.err
  show gpurun_out/bench_128_nb$nb.json
  SPCHOL_NB=$nb timeout 900 python bench.py --workload lap3d_64 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_64_nb$nb.json 2> gpurun_out/bench_64_nb$nb.err
  show gpurun_out/bench_64_nb$nb.json
done
SPCHOL_NB=512 SPCHOL_LOOKAHEAD=0 timeout 900 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_128_nb512_la0.json 2> gpurun_out/bench_128_nb512_la0.err
show gpurun_out/bench_128_nb512_la0.json
for f in gpurun_out/*.err; do tail -n 2 $f; done
