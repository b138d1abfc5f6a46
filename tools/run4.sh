#!/bin/bash
# session-13 first call: GPU tests, GEMM kernel microbenchmark (variants), 128^3 bench with K-class breakdown
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 600 tools/gemm_bench > gpurun_out/gemm_bench.jsonl 2>&1; echo "gemm_bench rc=$?"
timeout 300 tools/gemm_bench_pad4 v1 > gpurun_out/gemm_bench_pad4.jsonl 2>&1; echo "gemm_bench_pad4 rc=$?"
cat gpurun_out/gemm_bench.jsonl gpurun_out/gemm_bench_pad4.jsonl | cut -c1-200
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_128.json 2> gpurun_out/bench_128.err
echo "rc=$?"; cut -c1-3000 gpurun_out/bench_128.json; tail -5 gpurun_out/bench_128.err
