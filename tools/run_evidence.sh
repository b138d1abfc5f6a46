#!/bin/bash
# Round-2 evidence pass on ONE GPU (regenerates every single-GPU file DESIGN.md section 7 cites from the final code; ~6 GPU-minutes):
#   gpurun --timeout 2400 -- bash tools/run_evidence.sh
# default bench line (with `also`), reference arm, ncu launch list + --set full capture of the TMA update kernel at 128^3,
# direct lines of configs 3 and 5, 64^3, the update-kernel microbenchmark, and the GPU test suite.
mkdir -p gpurun_out
P=gpurun_out/r2z
timeout 1800 python -m pytest tests -q -m gpu > ${P}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 ${P}_pytest_gpu.log
timeout 1200 python bench.py > ${P}_bench_default.json 2> ${P}_bench_default.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > ${P}_bench_reference.json 2> ${P}_bench_reference.err; echo "ref rc=$?"
timeout 600 python bench.py --workload lap2d_1024 --steps 10 --warmup 3 > ${P}_bench_lap2d_1024.json 2> ${P}_2d.err; echo "2d rc=$?"
timeout 600 python bench.py --workload lap3d_64 --steps 10 --warmup 3 > ${P}_bench_lap3d_64.json 2> ${P}_64.err; echo "64 rc=$?"
timeout 600 python bench.py --workload batched_1138 --steps 5 --warmup 3 > ${P}_bench_batched_1138.json 2> ${P}_b.err; echo "batched rc=$?"
timeout 300 ./tools/gemm_bench > ${P}_gemm_microbench.jsonl 2> /dev/null; echo "microbench rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-also"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file ${P}_launches_lap3d_128.csv $CMD > ${P}_ncu_l128.log 2>&1
echo "launch list rc=$?"; wc -l ${P}_launches_lap3d_128.csv
timeout 1500 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_gemm_tma --csv --log-file ${P}_traffic_gemm_tma_lap3d_128.csv $CMD > ${P}_ncu_t128.log 2>&1
echo "traffic pass rc=$?"; wc -l ${P}_traffic_gemm_tma_lap3d_128.csv
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_gemm_tma -s 2400 -c 12 -f -o ${P}_ncu_gemm_tma_lap3d_128 $CMD > ${P}_ncu_f128.log 2>&1
echo "set full rc=$?"; ls -la ${P}_ncu_gemm_tma_lap3d_128.ncu-rep
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2z_bench_default.json').read().strip().splitlines()[-1])
print('128^3', round(d['ms_per_step'],1), 'e2e', round(d['e2e']['ms_per_step'],1), 'frac', d['roofline']['frac'], 'solve', d['solve']['seconds_host_to_host'], 'resid', d['residual']['rel_fro_estimate'])
for k,v in d.get('also',{}).items(): print(k, {kk: v.get(kk) for kk in ('ms_per_step','wall_seconds','same_input_speedup','error')}, (v.get('e2e') or {}).get('ms_per_step'))
PY
