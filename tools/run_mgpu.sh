#!/bin/bash
# N-GPU visit: NCCL parity worker, then the 128^3 bench with look-ahead on/off and a wider distributed block
mkdir -p gpurun_out
N=${1:-8}
export NCCL_DEBUG=WARN
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  tests/mgpu_worker.py > gpurun_out/mgpu_worker_$N.log 2>&1
echo "worker rc=$?" | tee -a gpurun_out/mgpu_worker_$N.log
tail -4 gpurun_out/mgpu_worker_$N.log
run() { name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 3 --warmup 3 --workload lap3d_128 --no-e2e $EXTRA > gpurun_out/${name}_n$N.json 2> gpurun_out/${name}_n$N.err
  echo "$name rc=$?"; python -c "
import json
for l in open('gpurun_out/${name}_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['ms_per_step'],1), round(d['value']), d['roofline']['whole_factor_frac_of_peak'])
        for r in d['per_rank']: print('   ', r['ms'], r['kernel_ms'])
"; tail -n 3 gpurun_out/${name}_n$N.err | cut -c1-300; }
EXTRA="" run b128 X=1
EXTRA="" run b128_la0 SPCHOL_LOOKAHEAD=0
EXTRA="--dist-nb 512" run b128_dnb512 X=1
