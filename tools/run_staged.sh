#!/bin/bash
# config 5 (batched) variant sweep for the next GPU session -- every variant below is validated on host threads
# (tests/test_path_emu.py) but only the default has been timed:   gpurun --timeout 900 -- bash tools/run_staged.sh
mkdir -p gpurun_out
B="python bench.py --workload batched_1138 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline"
run() { name=$1; shift; env "$@" timeout 120 $B > gpurun_out/batched_$name.json 2> gpurun_out/batched_$name.err
        echo "$name rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/batched_$name.json | head -1) $(grep -o '"max_rel_err_vs_oracle": [0-9.e-]*' gpurun_out/batched_$name.json)"; }
run default            SPCHOL_BATCH_PATH=1
run chain              SPCHOL_BATCH_PATH=0
for NU in 128 200 235; do run nu$NU SPCHOL_BATCH_PATH_NU=$NU; done
run subtree            SPCHOL_BATCH_GROUPS=1
run subtree_nu200      SPCHOL_BATCH_GROUPS=1 SPCHOL_BATCH_PATH_NU=200
for L0 in 2 1 0; do run subtree_level$L0 SPCHOL_BATCH_GROUPS=1 SPCHOL_BATCH_LEVEL0=$L0; done
# compile-time variants of k_front_path (csrc/path_kernel.cuh, PK_OPT bits: 1 corner size from the planner, 2 one warp
# barrier per pivot, 4 unmasked interior chunks, 8 L2 prefetch at group start), one at a time and all together
for OPT in 1 2 4 8 15; do
  SPCHOL_NVCC_DEFS="-DPK_OPT=$OPT" python -m sparsecholesky_b200.build --force > /dev/null && run opt$OPT SPCHOL_BATCH_PATH=1
done
SPCHOL_NVCC_DEFS="-DPK_OPT=15" python -m sparsecholesky_b200.build --force > /dev/null && run opt15_subtree SPCHOL_BATCH_GROUPS=1
python -m sparsecholesky_b200.build --force > /dev/null      # back to the shipped build
# read-back marks (OP_EXPORT, staged): bytes first, then the e2e number with and without
SPCHOL_TEST_STAGED=1 timeout 300 python -m pytest tests/test_chol_gpu.py -q -k "overlapped_readback" > gpurun_out/staged_marks.log 2>&1; echo "marks test rc=$?"; tail -n 3 gpurun_out/staged_marks.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_marks_off.json 2> gpurun_out/bench_marks_off.err; echo "marks off rc=$?"
SPCHOL_EXPORT_MARK_W=1024 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_marks_on.json 2> gpurun_out/bench_marks_on.err; echo "marks on rc=$?"
for f in gpurun_out/bench_marks_off.json gpurun_out/bench_marks_on.json; do echo "$f: $(grep -o '"e2e": {"value": [0-9.]*, "unit": "GFLOP/s", "ms_per_step": [0-9.]*' $f)"; done
