#!/bin/bash
# round 2: team kernel v6 (serial warp rotated over the teams of an SM, substitution instead of the 8x8 inverse), group cap sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_chol_gpu.py -q -m gpu -k "batched or subtree or golden or fullsize" > gpurun_out/r2l_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2l_pytest_gpu.log
B="python bench.py --workload batched_1138 --steps 5 --warmup 3 --no-cpu-baseline"
for C in "117 4" "117 2"; do set -- $C
  SPCHOL_BATCH_GROUP_M=$1 SPCHOL_BATCH_NW=$2 timeout 300 $B > gpurun_out/r2l_batched_g$1_nw$2.json 2> gpurun_out/r2l_batched_g$1_nw$2.err; echo "batched group=$1 nw=$2 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2l_batched_g$1_nw$2.json | head -2 | tr '\n' ' ') $(grep -o '"max_rel_err_vs_oracle": [0-9.e-]*' gpurun_out/r2l_batched_g$1_nw$2.json)"
done
tail -n 3 gpurun_out/r2l_batched_g117_nw4.err
for C in "64 2" "64 1" "48 1"; do set -- $C
  SPCHOL_TREE_CAP=$1 SPCHOL_TREE_NW=$2 timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2l_lap2d_1024_cap$1_nw$2.json 2> /dev/null; echo "2d cap=$1 nw=$2 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2l_lap2d_1024_cap$1_nw$2.json | head -1) $(grep -o '"small_fronts": [0-9.]*' gpurun_out/r2l_lap2d_1024_cap$1_nw$2.json)"
done
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-also > gpurun_out/r2l_lap3d_128_nomarks.json 2> gpurun_out/r2l_nomarks.err; echo "128 no marks rc=$? $(grep -o '"e2e": {"value": [0-9.]*, "unit": "GFLOP/s", "ms_per_step": [0-9.]*' gpurun_out/r2l_lap3d_128_nomarks.json) $(grep -o '"seconds_host_to_host": [0-9.]*' gpurun_out/r2l_lap3d_128_nomarks.json)"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tree_team -c 1 -f -o gpurun_out/r2l_ncu_team_2d \
  python bench.py --workload lap2d_1024 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2l_ncu2.log 2>&1; echo "ncu 2d rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tree_team -c 1 -f -o gpurun_out/r2l_ncu_team_batched \
  python bench.py --workload batched_1138 --nbatch 2048 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2l_ncu1.log 2>&1; echo "ncu batched rc=$?"
