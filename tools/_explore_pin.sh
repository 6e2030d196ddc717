#!/bin/bash
# scratch driver for one gpurun call: full GPU suite, smoke, default bench lines
mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/r2final_bench_p3m.json 2> gpurun_out/r2final_bench_p3m.err; tail -c 400 gpurun_out/r2final_bench_p3m.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2final_bench_p3m.json").read().strip().splitlines()[-1])
print("p3m", d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["e2e"].get("pipelined",{}).get("ms_per_step"), d["clocks"], d["gpu_launches"])
print(d["roofline"].get("exact_mode"), d["roofline"]["frac"], d["cpu_baseline"]["value"])
print({k:(v.get("ms_per_step"), v.get("value")) for k,v in d.get("secondary",{}).items()})
PY
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2final_bench_reference.json 2> gpurun_out/r2final_ref.err; tail -c 200 gpurun_out/r2final_bench_reference.json
