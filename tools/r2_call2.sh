#!/bin/bash
# round 2, second GPU visit: the warp-per-subtree kernel (k_tree_warp) -- parity, config 5 at several CTA shapes, config 3, ncu
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2b_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/r2b_pytest_gpu.log
B="python bench.py --workload batched_1138 --steps 5 --warmup 3 --no-cpu-baseline"
for W in 4 1 2 8; do
  SPCHOL_TREE_WPC=$W timeout 300 $B > gpurun_out/r2b_batched_wpc$W.json 2> gpurun_out/r2b_batched_wpc$W.err; echo "batched wpc=$W rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_batched_wpc$W.json | head -2 | tr '\n' ' ') $(grep -o '"max_rel_err_vs_oracle": [0-9.e-]*' gpurun_out/r2b_batched_wpc$W.json)"
done
tail -n 3 gpurun_out/r2b_batched_wpc4.err
for T in 1 0; do
  SPCHOL_TREE=$T timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2b_lap2d_1024_tree$T.json 2> gpurun_out/r2b_2d_tree$T.err; echo "2d tree=$T rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_lap2d_1024_tree$T.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2b_lap2d_1024_tree$T.json)"
done
for WK in 6e5 1e7; do
  SPCHOL_TREE_WORK=$WK timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2b_lap2d_1024_work$WK.json 2> /dev/null; echo "2d work=$WK $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_lap2d_1024_work$WK.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2b_lap2d_1024_work$WK.json)"
done
timeout 300 python bench.py --workload lap3d_64 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2b_lap3d_64.json 2> gpurun_out/r2b_64.err; echo "64 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_lap3d_64.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2b_lap3d_64.json)"
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2b_lap3d_128.json 2> gpurun_out/r2b_128.err; echo "128 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_lap3d_128.json | head -1) $(grep -o '"kernel_ms": {[^}]*}' gpurun_out/r2b_lap3d_128.json)"
# ncu: the batched kernel on 2048 matrices, and the panel kernels of the 2D problem
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tree_warp -c 1 -f -o gpurun_out/r2b_ncu_tree_batched \
  python bench.py --workload batched_1138 --nbatch 2048 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2b_ncu1.log 2>&1; echo "ncu batched rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_potrf_inv|k_tree_warp|k_trsm_inv" -s 40 -c 6 -f -o gpurun_out/r2b_ncu_2d_panel \
  python bench.py --workload lap2d_1024 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2b_ncu2.log 2>&1; echo "ncu 2d rc=$?"
ls -la gpurun_out/*.ncu-rep
