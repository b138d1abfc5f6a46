#!/bin/bash
# round 2: team kernel v5 (look-ahead micro-panels), solve with resident buffers, read-back marks, default bench line with `also`
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_chol_gpu.py tests/test_dropin_cpp.py -q -m gpu > gpurun_out/r2j_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r2j_pytest_gpu.log
B="python bench.py --workload batched_1138 --steps 5 --warmup 3 --no-cpu-baseline"
for C in "117 4" "117 2"; do set -- $C
  SPCHOL_BATCH_GROUP_M=$1 SPCHOL_BATCH_NW=$2 timeout 300 $B > gpurun_out/r2j_batched_g$1_nw$2.json 2> gpurun_out/r2j_batched_g$1_nw$2.err; echo "batched group=$1 nw=$2 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2j_batched_g$1_nw$2.json | head -2 | tr '\n' ' ') $(grep -o '"max_rel_err_vs_oracle": [0-9.e-]*' gpurun_out/r2j_batched_g$1_nw$2.json)"
done
tail -n 3 gpurun_out/r2j_batched_g117_nw4.err
for C in "64 2" "64 1" "48 1" "48 2"; do set -- $C
  SPCHOL_TREE_CAP=$1 SPCHOL_TREE_NW=$2 timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2j_lap2d_1024_cap$1_nw$2.json 2> /dev/null; echo "2d cap=$1 nw=$2 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2j_lap2d_1024_cap$1_nw$2.json | head -1) $(grep -o '"small_fronts": [0-9.]*' gpurun_out/r2j_lap2d_1024_cap$1_nw$2.json)"
done
SPCHOL_EXPORT_MARK_W=1024 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-also > gpurun_out/r2j_lap3d_128_marks.json 2> gpurun_out/r2j_marks.err; echo "128 marks rc=$? $(grep -o '"e2e": {"value": [0-9.]*, "unit": "GFLOP/s", "ms_per_step": [0-9.]*' gpurun_out/r2j_lap3d_128_marks.json) $(grep -o '"mismatching_chunks": [0-9]*' gpurun_out/r2j_lap3d_128_marks.json)"
/usr/bin/time -v timeout 900 python bench.py > gpurun_out/r2j_bench_default.json 2> gpurun_out/r2j_bench_default.err; echo "default rc=$?"; grep -E "Elapsed|Maximum resident" gpurun_out/r2j_bench_default.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2j_bench_default.json').read().strip().splitlines()[-1])
print('128^3', round(d['ms_per_step'],1), 'e2e', round(d['e2e']['ms_per_step'],1), 'frac', d['roofline']['frac'], 'solve', d['solve'], 'resid', d['residual']['rel_fro_estimate'])
for k,v in d.get('also',{}).items(): print(k, {kk: v.get(kk) for kk in ('ms_per_step','wall_seconds','same_input_speedup','error')}, (v.get('e2e') or {}).get('ms_per_step'))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tree_team -c 1 -f -o gpurun_out/r2j_ncu_team_batched \
  python bench.py --workload batched_1138 --nbatch 2048 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2j_ncu1.log 2>&1; echo "ncu batched rc=$?"
