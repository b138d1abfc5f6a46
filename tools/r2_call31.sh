#!/bin/bash
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
timeout 80 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_worker.py > gpurun_out/r2z_mgpu_worker_n2.log 2>&1
echo "worker rc=$?"; grep "mgpu\|MGPU\|Error" gpurun_out/r2z_mgpu_worker_n2.log | cut -c1-200
