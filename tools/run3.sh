#!/bin/bash
mkdir -p gpurun_out
python bench.py --workload lap3d_64 --steps 5 --warmup 3 > gpurun_out/bench_64.json 2> gpurun_out/bench_64.err
echo "rc=$?"; cat gpurun_out/bench_64.json; tail -5 gpurun_out/bench_64.err
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_128.json 2> gpurun_out/bench_128.err
echo "rc=$?"; cat gpurun_out/bench_128.json; tail -5 gpurun_out/bench_128.err
python bench.py --workload lap3d_64 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 2163 -c 730 --csv --log-file gpurun_out/launches_64.csv \
  python bench.py --workload lap3d_64 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
