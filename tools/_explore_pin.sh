# scratch driver for gpurun calls (development only)
set -x
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -25 > gpurun_out/r2l_gputests.log; tail -12 gpurun_out/r2l_gputests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu > gpurun_out/r2l_bench_p3m.json 2> gpurun_out/r2l_bench_p3m.err; tail -c 300 gpurun_out/r2l_bench_p3m.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2l_bench_p3m.json").read().strip().splitlines()[-1]); print(d["ms_per_step"], d["phase_ms"], d["e2e"]["ms_per_step"])
PY
