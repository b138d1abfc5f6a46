#!/bin/bash
# round 2, third GPU visit: the team kernel (k_tree_team: NW warps per matrix / subtree, DMMA update)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_chol_gpu.py -x -q -m gpu > gpurun_out/r2d_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r2d_pytest_gpu.log
B="python bench.py --workload batched_1138 --steps 5 --warmup 3 --no-cpu-baseline"
for W in 4 2 1; do
  SPCHOL_BATCH_NW=$W timeout 300 $B > gpurun_out/r2d_batched_nw$W.json 2> gpurun_out/r2d_batched_nw$W.err; echo "batched nw=$W rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2d_batched_nw$W.json | head -2 | tr '\n' ' ') $(grep -o '"max_rel_err_vs_oracle": [0-9.e-]*' gpurun_out/r2d_batched_nw$W.json)"
done
tail -n 3 gpurun_out/r2d_batched_nw4.err
for W in 1 2 4; do
  SPCHOL_TREE_NW=$W timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2d_lap2d_1024_nw$W.json 2> gpurun_out/r2d_2d_nw$W.err; echo "2d nw=$W rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2d_lap2d_1024_nw$W.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2d_lap2d_1024_nw$W.json)"
done
for CAP in 64 32; do
  SPCHOL_TREE_CAP=$CAP SPCHOL_TREE_NW=1 timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2d_lap2d_1024_cap$CAP.json 2> /dev/null; echo "2d cap=$CAP $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2d_lap2d_1024_cap$CAP.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2d_lap2d_1024_cap$CAP.json)"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tree_team -c 1 -f -o gpurun_out/r2d_ncu_team_batched \
  python bench.py --workload batched_1138 --nbatch 2048 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2d_ncu1.log 2>&1; echo "ncu batched rc=$?"
