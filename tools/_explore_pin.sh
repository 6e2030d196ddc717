#!/bin/bash
# scratch driver for one gpurun call: N-GPU bench with / without host binding
N=${1:-4}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2v_topo_${N}gpu.txt 2>&1; nproc >> gpurun_out/r2v_topo_${N}gpu.txt; lscpu | grep -i "numa\|socket\|model name" >> gpurun_out/r2v_topo_${N}gpu.txt
for b in ${2:-1 0}; do
NB_BIND_HOST=$b timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2v_bench_p3m_${N}gpu_bind$b.json 2> gpurun_out/r2v_bench_p3m_${N}gpu_bind$b.err
echo "rc=$?"; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2v_bench_p3m_${N}gpu_bind$b.err | tail -c 800
python - <<PY
import json
d=json.loads(open("gpurun_out/r2v_bench_p3m_${N}gpu_bind$b.json").read().strip().splitlines()[-1])
print("N=$N bind=$b", round(d["ms_per_step"],2), "e2e ms", round(d["e2e"]["ms_per_step"],2), d["e2e"].get("host_binding_rank0"))
PY
done
cat gpurun_out/r2v_topo_${N}gpu.txt | head -30
