#!/bin/bash
# the last GPU seconds of round 1: config 5 with the final path-group kernel (the bench checks three matrices against the oracle)
mkdir -p gpurun_out
timeout 60 python bench.py --workload batched_1138 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/batched_final.json 2> gpurun_out/batched_final.err; echo "rc=$?"
grep -o '"ms_per_step": [0-9.]*' gpurun_out/batched_final.json | head -1; grep -o '"parity": {[^}]*}' gpurun_out/batched_final.json; tail -n 2 gpurun_out/batched_final.err
