#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
export NCCL_DEBUG=WARN
timeout 300 python -m pytest tests/test_chol_gpu.py -q -m gpu -k "solve" 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  tests/mgpu_worker.py > gpurun_out/r2r_mgpu_worker_$N.log 2>&1
echo "worker rc=$?" | tee -a gpurun_out/r2r_mgpu_worker_$N.log
grep "mgpu\|MGPU\|Error" gpurun_out/r2r_mgpu_worker_$N.log | cut -c1-300 | head -20
