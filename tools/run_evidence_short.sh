#!/bin/bash
# Trimmed evidence pass on ONE GPU after the last kernel change (~9 GPU-minutes): GPU tests, default bench line, configs 3 / 5 / 64^3,
# update-kernel microbenchmark, ncu launch list and one --set full capture of k_gemm_tma at 128^3 (1 warm-up step: the capture
# window -s 732 is launch 176 of the second factorization, the same place as -s 2400 with 3 warm-up steps in run_evidence.sh)
mkdir -p gpurun_out
P=gpurun_out/r2y
timeout 900 python -m pytest tests -q -m gpu > ${P}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 2 ${P}_pytest_gpu.log
timeout 600 python bench.py > ${P}_bench_default.json 2> ${P}_bench_default.err; echo "bench rc=$?"
timeout 300 python bench.py --workload lap2d_1024 --steps 10 --warmup 3 --no-also > ${P}_bench_lap2d_1024.json 2> ${P}_2d.err; echo "2d rc=$?"
timeout 300 python bench.py --workload lap3d_64 --steps 10 --warmup 3 --no-also > ${P}_bench_lap3d_64.json 2> ${P}_64.err; echo "64 rc=$?"
timeout 300 python bench.py --workload batched_1138 --steps 5 --warmup 3 --no-also > ${P}_bench_batched_1138.json 2> ${P}_b.err; echo "batched rc=$?"
timeout 200 ./tools/gemm_bench > ${P}_gemm_microbench.jsonl 2> /dev/null; echo "microbench rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-also"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file ${P}_launches_lap3d_128.csv $CMD > ${P}_ncu_l128.log 2>&1
echo "launch list rc=$?"; wc -l ${P}_launches_lap3d_128.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gemm_tma -s 732 -c 12 -f -o ${P}_ncu_gemm_tma_lap3d_128 $CMD > ${P}_ncu_f128.log 2>&1
echo "set full rc=$?"; ls -la ${P}_ncu_gemm_tma_lap3d_128.ncu-rep
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2y_bench_default.json').read().strip().splitlines()[-1])
print('128^3', round(d['ms_per_step'],1), 'e2e', round(d['e2e']['ms_per_step'],1), 'frac', d['roofline']['frac'], 'solve', d['solve']['seconds_host_to_host'], 'resid', d['residual']['rel_fro_estimate'])
for k,v in d.get('also',{}).items(): print(k, {kk: v.get(kk) for kk in ('ms_per_step','wall_seconds','same_input_speedup','error')}, (v.get('e2e') or {}).get('ms_per_step'))
PY
