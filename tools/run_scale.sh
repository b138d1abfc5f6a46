#!/bin/bash
# N-GPU bench lines of the headline workload on the final code (run once per N: gpurun --gpus N -- bash tools/run_scale.sh N)
mkdir -p gpurun_out
N=${1:-8}
export NCCL_DEBUG=WARN
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 5 --warmup 3 --workload lap3d_128 > gpurun_out/r2z_bench_lap3d_128_n$N.json 2> gpurun_out/r2z_bench_n$N.err
echo "bench rc=$?"; python - <<PY
import json
for l in open('gpurun_out/r2z_bench_lap3d_128_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l); print('N=$N', round(d['ms_per_step'],1), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],1), 'solve', d['solve']['seconds_host_to_host'], d['solve']['rel_residual'], 'd2h GB', round(d['e2e']['d2h_bytes_per_step']/1e9,1), 'residual', d['residual']['rel_fro_estimate'])
        for r in d['per_rank']: print('   ', r['ms'], r['kernel_ms'])
PY
tail -n 2 gpurun_out/r2z_bench_n$N.err | cut -c1-300
if [ "$N" = "8" ]; then
  timeout 90 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus 8 --steps 3 --warmup 3 --workload batched_1138 > gpurun_out/r2z_bench_batched_1138_n8.json 2> gpurun_out/r2z_bench_batched_n8.err
  echo "batched n8 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2z_bench_batched_1138_n8.json | head -2 | tr '\n' ' ')"
  timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 \
    tests/mgpu_worker.py > gpurun_out/r2z_mgpu_worker_n8.log 2>&1
  echo "worker rc=$?"; grep "MGPU OK" gpurun_out/r2z_mgpu_worker_n8.log
fi
