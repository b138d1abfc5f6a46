#!/bin/bash
# round 2: TMA-staged DMMA update kernel -- microbenchmark against the round-1 kernel, parity suite, 128^3 / 64^3 / 2D benches on/off
mkdir -p gpurun_out
timeout 600 ./tools/gemm_bench > gpurun_out/r2g_gemm_microbench.jsonl 2> gpurun_out/r2g_gemm_microbench.err; echo "microbench rc=$?"; cat gpurun_out/r2g_gemm_microbench.jsonl | cut -c1-220
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/r2g_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r2g_pytest_gpu.log
for T in 1 0; do
  SPCHOL_GEMM_TMA=$T timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-also > gpurun_out/r2g_lap3d_128_tma$T.json 2> gpurun_out/r2g_128_tma$T.err; echo "128 tma=$T rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2g_lap3d_128_tma$T.json | head -1) $(grep -o '"by_K_class": {.*}}, "whole' gpurun_out/r2g_lap3d_128_tma$T.json)"
  SPCHOL_GEMM_TMA=$T timeout 300 python bench.py --workload lap3d_64 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2g_lap3d_64_tma$T.json 2> /dev/null; echo "64 tma=$T $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2g_lap3d_64_tma$T.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2g_lap3d_64_tma$T.json)"
  SPCHOL_GEMM_TMA=$T timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2g_lap2d_1024_tma$T.json 2> /dev/null; echo "2d tma=$T $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2g_lap2d_1024_tma$T.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2g_lap2d_1024_tma$T.json)"
done
tail -n 3 gpurun_out/r2g_128_tma1.err
for C in "64 2" "64 4" "96 2" "48 1"; do set -- $C
  SPCHOL_TREE_CAP=$1 SPCHOL_TREE_NW=$2 timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2g_lap2d_1024_cap$1_nw$2.json 2> /dev/null; echo "2d cap=$1 nw=$2 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2g_lap2d_1024_cap$1_nw$2.json | head -1) $(grep -o '"small_fronts": [0-9.]*' gpurun_out/r2g_lap2d_1024_cap$1_nw$2.json)"
done
