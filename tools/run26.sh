#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -8 gpurun_out/pytest_gpu.log | cut -c1-220
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
