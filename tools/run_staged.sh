#!/bin/bash
# first GPU call of the next session: does the staged path-group kernel (SPCHOL_BATCH_PATH=1) work on the device,
# and what does it buy on config 5?   gpurun --timeout 900 -- bash tools/run_staged.sh
mkdir -p gpurun_out
SPCHOL_TEST_STAGED=1 timeout 600 python -m pytest tests/test_chol_gpu.py -q -k "path_groups or batched" > gpurun_out/staged_tests.log 2>&1
echo "staged tests rc=$?"; tail -n 5 gpurun_out/staged_tests.log
B="python bench.py --workload batched_1138 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 600 $B > gpurun_out/batched_chain.json 2> gpurun_out/batched_chain.err; echo "chain rc=$?"
for NU in 128 160 200; do
  SPCHOL_BATCH_PATH=1 SPCHOL_BATCH_PATH_NU=$NU timeout 600 $B > gpurun_out/batched_path_$NU.json 2> gpurun_out/batched_path_$NU.err; echo "path nu=$NU rc=$?"
done
for f in gpurun_out/batched_chain.json gpurun_out/batched_path_*.json; do echo "$f: $(grep -o '"ms_per_step": [0-9.]*' $f | head -1)"; done
# where the time goes inside the new kernel: one capture of k_front_path (a 2048-matrix batch keeps the replay short)
SPCHOL_BATCH_PATH=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_front_path -c 1 -f -o gpurun_out/prof_front_path \
  python bench.py --workload batched_1138 --nbatch 2048 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_front_path.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_front_path.ncu-rep 2>/dev/null
