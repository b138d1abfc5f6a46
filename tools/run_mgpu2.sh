#!/bin/bash
# 2-GPU visit: NCCL parity worker, then the bench at 64^3 and 128^3 on 2 ranks
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/gpus.txt 2>&1
N=${1:-2}
export NCCL_DEBUG=WARN
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  tests/mgpu_worker.py > gpurun_out/mgpu_worker_$N.log 2>&1
echo "worker rc=$?" | tee -a gpurun_out/mgpu_worker_$N.log
tail -15 gpurun_out/mgpu_worker_$N.log
for wl in lap3d_64 lap3d_128; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 2 --warmup 3 --workload $wl --no-e2e > gpurun_out/bench_${wl}_n$N.json 2> gpurun_out/bench_${wl}_n$N.err
  echo "bench $wl rc=$?"; cat gpurun_out/bench_${wl}_n$N.json; tail -5 gpurun_out/bench_${wl}_n$N.err
done
