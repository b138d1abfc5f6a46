#!/bin/bash
mkdir -p gpurun_out
N=2
export NCCL_DEBUG=WARN
SPCHOL_TRACE=gpurun_out/trace128n2.txt timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 1 --warmup 3 --workload lap3d_128 --no-e2e > gpurun_out/tr_n$N.json 2> gpurun_out/tr_n$N.err
echo rc=$?; cut -c1-400 gpurun_out/tr_n$N.json; tail -n 3 gpurun_out/tr_n$N.err | cut -c1-300
ls -la gpurun_out/trace128n2*
