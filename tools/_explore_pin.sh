#!/bin/bash
# scratch driver for one gpurun call: presort overlap sweep
mkdir -p gpurun_out
python -m pytest tests/test_gpu_force.py -q -m gpu -x 2>&1 | tail -2
for cfg in "0 88" "1 88" "1 96" "1 104"; do
  set -- $cfg
  NB_P3M_PRESORT=$1 NB_P3M_SMS=$2 python bench.py --steps 6 --warmup 3 --no-secondary --no-cpu > gpurun_out/r2w_bench_presort$1_sms$2.json 2> gpurun_out/r2w.err || tail -3 gpurun_out/r2w.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2w_bench_presort$1_sms$2.json").read().strip().splitlines()[-1])
print("presort $1 sms $2:", round(d["ms_per_step"],2), round(d["e2e"]["ms_per_step"],1), d["clocks"]["power_w_max"])
PY
done
