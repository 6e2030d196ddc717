#!/bin/bash
mkdir -p gpurun_out
for rn in 20 12 8; do
  NB_P3M_RNEAR=$rn python -m pytest tests/test_gpu_baseline_sizes.py -q -m gpu -x -s -k "exact" 2>&1 | grep -i "P3M 128\|near pairs\|failed\|rror" | head -3
done
for cfg in "1 20" "1 12"; do
  set -- $cfg
  NB_P3M_NEAR=$1 NB_P3M_RNEAR=$2 python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu > gpurun_out/r2z_bench_near$1_rn$2.json 2> gpurun_out/r2z.err || tail -3 gpurun_out/r2z.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2z_bench_near$1_rn$2.json").read().strip().splitlines()[-1])
print("near $1 rnear $2:", round(d["ms_per_step"],2), round(d["phase_ms"]["p3m_short"],2))
PY
done
