#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
export NCCL_DEBUG=WARN
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  tests/mgpu_worker.py > gpurun_out/mgpu_worker_$N.log 2>&1
echo "worker rc=$?" | tee -a gpurun_out/mgpu_worker_$N.log
tail -3 gpurun_out/mgpu_worker_$N.log
bash tools/run_scale.sh $N
