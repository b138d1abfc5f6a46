#!/bin/bash
# the last GPU minutes of a round: staged path-group kernel, correctness first, then one timing pair
mkdir -p gpurun_out
SPCHOL_TEST_STAGED=1 timeout 90 python -m pytest tests/test_chol_gpu.py -q -x -k "path_groups" > gpurun_out/staged_tests.log 2>&1
echo "staged tests rc=$?"; tail -n 4 gpurun_out/staged_tests.log
B="python bench.py --workload batched_1138 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
SPCHOL_BATCH_PATH=1 timeout 100 $B > gpurun_out/batched_path_160.json 2> gpurun_out/batched_path_160.err; echo "path rc=$?"
grep -o '"ms_per_step": [0-9.]*' gpurun_out/batched_path_160.json | head -1
timeout 100 $B > gpurun_out/batched_chain.json 2> gpurun_out/batched_chain.err; echo "chain rc=$?"
grep -o '"ms_per_step": [0-9.]*' gpurun_out/batched_chain.json | head -1
