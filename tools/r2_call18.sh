#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/r2n_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2n_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2n_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/r2n_smoke.log
S=$(date +%s); timeout 1200 python bench.py > gpurun_out/r2n_bench_default.json 2> gpurun_out/r2n_bench_default.err; echo "default rc=$? wall $(( $(date +%s) - S )) s"
S=$(date +%s); timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r2n_bench_reference.json 2> /dev/null; echo "reference rc=$? wall $(( $(date +%s) - S )) s"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2n_bench_default.json').read().strip().splitlines()[-1])
print('128^3', round(d['ms_per_step'],1), 'e2e', round(d['e2e']['ms_per_step'],1), 'frac', d['roofline']['frac'], 'traffic', d['roofline']['traffic'], 'solve', d['solve']['seconds_host_to_host'], 'resid', d['residual']['rel_fro_estimate'])
print(d['kernel_ms'])
for k,v in d.get('also',{}).items(): print(k, {kk: v.get(kk) for kk in ('ms_per_step','wall_seconds','same_input_speedup','error')}, (v.get('e2e') or {}).get('ms_per_step'))
PY
timeout 300 python bench.py --workload lap2d_1024 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2n_lap2d_1024.json 2> /dev/null; echo "2d $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2n_lap2d_1024.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2n_lap2d_1024.json) $(grep -o '"seconds_host_to_host": [0-9.]*' gpurun_out/r2n_lap2d_1024.json)"
