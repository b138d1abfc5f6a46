#!/bin/bash
mkdir -p gpurun_out
timeout 280 python bench.py > gpurun_out/bench_final_n1.json 2> gpurun_out/bench_final_n1.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/bench_final_n1.json
