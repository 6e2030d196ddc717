#!/bin/bash
# scratch driver for one gpurun call: near-pair list accuracy and cost
mkdir -p gpurun_out
for nr in 0 1; do
  NB_P3M_NEAR=$nr python -m pytest tests/test_gpu_baseline_sizes.py -q -m gpu -s -k "p3m_128 or c5_p3m" 2>&1 | grep -i "P3M 128\|C5 short\|C5 P3M total\|passed\|failed\|Error\|error" | head -12
done
python -m pytest tests/test_gpu_force.py tests/test_gpu_dist.py -q -m gpu -x 2>&1 | tail -3
for nr in 0 1; do
  NB_P3M_NEAR=$nr python bench.py --steps 6 --warmup 3 --no-secondary --no-cpu > gpurun_out/r2x_bench_near$nr.json 2> gpurun_out/r2x.err || tail -3 gpurun_out/r2x.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2x_bench_near$nr.json").read().strip().splitlines()[-1])
print("near $nr:", round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms"].items()}, round(d["e2e"]["ms_per_step"],1))
PY
done
