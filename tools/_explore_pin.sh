#!/bin/bash
# scratch driver for one gpurun call: N-GPU bench (N from $1), overlapped (default)
N=${1:-2}
mkdir -p gpurun_out
if [ "$2" = "tests" ]; then python -m pytest tests/test_gpu_nccl.py tests/test_gpu_force.py -q -m gpu -x 2>&1 | tail -3; fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2s_bench_p3m_${N}gpu.json 2> gpurun_out/r2s_bench_p3m_${N}gpu.err
echo "rc=$?"; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2s_bench_p3m_${N}gpu.err | tail -c 1500
python - <<PY
import json
d=json.loads(open("gpurun_out/r2s_bench_p3m_${N}gpu.json").read().strip().splitlines()[-1])
print("N=$N", d["ms_per_step"], d["value"], d["e2e"]["value"] if d.get("e2e") else None)
print(d["phase_ms"]); print(d["roofline"]["kernel"][:20], d["roofline"]["ms"], d["roofline"]["frac"])
print(d["invariants"]["mom_sq_sum"], d["invariants"]["pos_sum"])
PY
