#!/bin/bash
# final round-1 numbers on N GPUs: 128^3 (and the batched config) through bench.py exactly as the driver launches it
mkdir -p gpurun_out
N=${1:-8}
export NCCL_DEBUG=WARN
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/final_lap3d_128_n$N.json 2> gpurun_out/final_lap3d_128_n$N.err
echo "128 rc=$?"; python -c "
import json
for l in open('gpurun_out/final_lap3d_128_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l); print('n=$N', round(d['ms_per_step'],1), round(d['value']), d['roofline']['whole_factor_frac_of_peak'], 'e2e', d['e2e'] and round(d['e2e']['ms_per_step'],1))
        for r in d.get('per_rank',[]): print('   ', r['ms'], r['kernel_ms'])
"; tail -n 2 gpurun_out/final_lap3d_128_n$N.err | cut -c1-300
if [ "$2" = "batched" ]; then
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --steps 3 --warmup 3 --workload batched_1138 > gpurun_out/final_batched_n$N.json 2> gpurun_out/final_batched_n$N.err
echo "batched rc=$?"; grep '^{' gpurun_out/final_batched_n$N.json | cut -c1-400; tail -n 2 gpurun_out/final_batched_n$N.err | cut -c1-300
fi
