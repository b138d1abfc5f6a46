#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 900 python tools/quick_time.py 1024x1024x1 64x64x64 96x96x96 > gpurun_out/quick_time.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/quick_time.log
