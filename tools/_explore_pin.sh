#!/bin/bash
# scratch driver for one gpurun call (development only): the other three workloads
mkdir -p gpurun_out
for w in pk pm pp; do
  python bench.py --workload $w --steps 3 --warmup 3 --no-cpu > gpurun_out/r2final_bench_$w.json 2> gpurun_out/r2final_$w.err || tail -c 600 gpurun_out/r2final_$w.err
  python -c "
import json
d=json.loads(open('gpurun_out/r2final_bench_$w.json').read().strip().splitlines()[-1])
print('$w', d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['roofline']['kernel'][:30], round(d['roofline']['frac'],3))"
done
