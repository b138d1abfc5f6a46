#!/bin/bash
mkdir -p gpurun_out
SPCHOL_TRACE=gpurun_out/trace128.txt SPCHOL_WIDE_W=1 timeout 900 python bench.py --workload lap3d_128 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/tr128.json 2> gpurun_out/tr128.err
tail -n 2 gpurun_out/tr128.err; wc -l gpurun_out/trace128.txt
