#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --workload batched_1138 --steps 3 --warmup 3 > gpurun_out/bench_batched_n1.json 2> gpurun_out/bench_batched_n1.err
echo "rc=$?"; cut -c1-1800 gpurun_out/bench_batched_n1.json; tail -5 gpurun_out/bench_batched_n1.err
