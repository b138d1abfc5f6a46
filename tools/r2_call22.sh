#!/bin/bash
# distributed solve: NCCL parity worker (factor + residual + solve on every rank), then the bench at N ranks with the solve
mkdir -p gpurun_out
N=${1:-2}
export NCCL_DEBUG=WARN
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  tests/mgpu_worker.py > gpurun_out/r2r_mgpu_worker_$N.log 2>&1
echo "worker rc=$?" | tee -a gpurun_out/r2r_mgpu_worker_$N.log
tail -12 gpurun_out/r2r_mgpu_worker_$N.log | cut -c1-400
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 3 --warmup 3 --workload lap3d_128 --no-also > gpurun_out/r2r_b128_n$N.json 2> gpurun_out/r2r_b128_n$N.err
echo "bench rc=$?"; tail -3 gpurun_out/r2r_b128_n$N.err | cut -c1-300
python - <<PY
import json
for l in open('gpurun_out/r2r_b128_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l); print(round(d['ms_per_step'],1), 'e2e', d['e2e'], 'solve', d['solve'], 'resid', d['residual']['rel_fro_estimate'])
PY
