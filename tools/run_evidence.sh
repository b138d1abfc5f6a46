#!/bin/bash
# round-1 evidence run on 1 GPU: default bench line, ncu launch list of the same command, --set full capture of the DMMA kernel
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
cut -c1-600 gpurun_out/bench_default.json; tail -n 2 gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
cut -c1-300 gpurun_out/bench_reference.json
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s 12600 -c 4300 --csv --log-file gpurun_out/launches_128.csv $CMD > gpurun_out/ncu_l128.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_128.csv
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_gemm_dmma -s 2455 -c 12 -f -o gpurun_out/prof_gemm_128 $CMD > gpurun_out/ncu_f128.log 2>&1
echo "set full rc=$?"; ls -la gpurun_out/prof_gemm_128.ncu-rep
timeout 600 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 > gpurun_out/bench_2d.json 2> gpurun_out/bench_2d.err; echo "2d rc=$?"
timeout 600 python bench.py --workload lap3d_64 --steps 5 --warmup 3 > gpurun_out/bench_64.json 2> gpurun_out/bench_64.err; echo "64 rc=$?"
timeout 600 python bench.py --workload batched_1138 --steps 3 --warmup 3 > gpurun_out/bench_batched.json 2> gpurun_out/bench_batched.err; echo "batched rc=$?"
