#!/bin/bash
# first GPU visit: peaks, parity tests, a quick timing of the working path
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
./tools/fp64_peak > gpurun_out/fp64_peak.jsonl 2>&1
python tools/dgemm_peak.py >> gpurun_out/fp64_peak.jsonl 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/quick_time.py > gpurun_out/quick_time.log 2>&1
tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/fp64_peak.jsonl; tail -30 gpurun_out/quick_time.log
