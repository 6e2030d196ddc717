#!/bin/bash
# scratch driver for one gpurun call (development only): N-GPU default bench (N from $1)
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/last_bench_${N}gpu.json 2> gpurun_out/last_bench_${N}gpu.err
echo "rc=$?"; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/last_bench_${N}gpu.err | tail -c 600
python -c "
import json
d=json.loads(open('gpurun_out/last_bench_${N}gpu.json').read().strip().splitlines()[-1])
print('N=$N', round(d['ms_per_step'],2), d['value'], round(d['e2e']['ms_per_step'],1), d['invariants']['n_particles_total'], d['invariants']['mom_sq_sum'])"
