# scratch driver for gpurun calls (development only)
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2e_gputests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2e_gputests.log; tail -8 gpurun_out/r2e_gputests.log
python tools/pin_concept.py --backend gpu --seeds 4 --methods pm,p3m,snapvspk --lattice-offset 0.0 --out gpurun_out/r2e_pin_off0.json > gpurun_out/r2e_pin_off0.log 2>&1
tail -12 gpurun_out/r2e_pin_off0.log
python tools/pin_concept.py --backend gpu --seeds 8 --methods giant32 --lattice-offset 0.0 --out gpurun_out/r2e_pin_giant32_off0.json > gpurun_out/r2e_pin_giant32_off0.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 8 --methods giant32 --lattice-offset 0.5 --out gpurun_out/r2e_pin_giant32_off05.json > gpurun_out/r2e_pin_giant32_off05.log 2>&1
tail -4 gpurun_out/r2e_pin_giant32*.log
