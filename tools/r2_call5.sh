#!/bin/bash
# round 2: N-GPU visit -- NCCL parity worker on the final code (incl. the distributed residual), then the 128^3 bench
mkdir -p gpurun_out
N=${1:-2}
export NCCL_DEBUG=WARN
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  tests/mgpu_worker.py > gpurun_out/r2h_mgpu_worker_n$N.log 2>&1
echo "worker rc=$?" | tee -a gpurun_out/r2h_mgpu_worker_n$N.log
tail -6 gpurun_out/r2h_mgpu_worker_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 3 --warmup 3 --workload lap3d_128 > gpurun_out/r2h_bench_lap3d_128_n$N.json 2> gpurun_out/r2h_bench_n$N.err
echo "bench rc=$?"; python - <<PY
import json
for l in open('gpurun_out/r2h_bench_lap3d_128_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l); print('N=$N', round(d['ms_per_step'],1), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],1), 'd2h GB', round(d['e2e']['d2h_bytes_per_step']/1e9,1), 'residual', d['residual'])
        for r in d['per_rank']: print('   ', r['ms'], r['kernel_ms'])
PY
tail -n 3 gpurun_out/r2h_bench_n$N.err | cut -c1-300
