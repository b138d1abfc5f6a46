#!/bin/bash
# round 2, first GPU visit: GPU parity suite, then baseline numbers of every bench workload on the round-1 kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2a_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/r2a_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2a_bench_lap3d_128.json 2> gpurun_out/r2a_bench_lap3d_128.err; echo "128 rc=$?"; cut -c1-400 gpurun_out/r2a_bench_lap3d_128.json
timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 > gpurun_out/r2a_bench_lap2d_1024.json 2> gpurun_out/r2a_2d.err; echo "2d rc=$?"; cut -c1-300 gpurun_out/r2a_bench_lap2d_1024.json
timeout 300 python bench.py --workload batched_1138 --steps 3 --warmup 3 > gpurun_out/r2a_bench_batched.json 2> gpurun_out/r2a_b.err; echo "batched rc=$?"; cut -c1-300 gpurun_out/r2a_bench_batched.json
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,memory.total --format=csv
