#!/bin/bash
# scratch driver for one gpurun call: N-GPU default bench (N from $1)
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2final_bench_p3m_${N}gpu.json 2> gpurun_out/r2final_bench_p3m_${N}gpu.err
echo "rc=$?"; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2final_bench_p3m_${N}gpu.err | tail -c 1000
python - <<PY
import json
d=json.loads(open("gpurun_out/r2final_bench_p3m_${N}gpu.json").read().strip().splitlines()[-1])
print("N=$N", round(d["ms_per_step"],2), d["value"], round(d["e2e"]["ms_per_step"],1), d["clocks"])
print(d["phase_ms"]); print(d["invariants"]["n_particles_total"], d["invariants"]["mom_sq_sum"])
PY
