#!/bin/bash
# look-ahead A/B on one GPU + GPU test-suite
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
for la in 1 0; do
  SPCHOL_LOOKAHEAD=$la timeout 600 python bench.py --workload lap3d_64 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_64_la$la.json 2> gpurun_out/bench_64_la$la.err
  echo "64 la=$la rc=$?"; python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_64_la$la.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['kernel_ms'], d['roofline']['by_K_class'])"
  SPCHOL_LOOKAHEAD=$la timeout 900 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_128_la$la.json 2> gpurun_out/bench_128_la$la.err
  echo "128 la=$la rc=$?"; python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_128_la$la.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['kernel_ms'], d['roofline']['by_K_class'])"
done
for nb in 512; do
  SPCHOL_NB=$nb timeout 900 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_128_nb$nb.json 2> gpurun_out/bench_128_nb$nb.err
  echo "128 NB=$nb rc=$?"; python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_128_nb$nb.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['kernel_ms'], d['roofline']['by_K_class'])"
done
timeout 600 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_2d.json 2> gpurun_out/bench_2d.err
python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_2d.json').read().strip().splitlines()[-1])
print('2d', d['ms_per_step'], d['value'], d['kernel_ms'])"
tail -3 gpurun_out/*.err
