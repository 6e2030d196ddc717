set -x
python tools/pin_concept.py --backend gpu --seeds 4 --methods pm,p3m,snapvspk --lattice-offset 0.5 --estimator-interlace 0 --out gpurun_out/r2c_pin_sub_off0.5_il0.json > gpurun_out/r2c_pin_sub_off0.5_il0.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 4 --methods pm,p3m,snapvspk --lattice-offset 0.5 --estimator-interlace 1 --out gpurun_out/r2c_pin_sub_off0.5_il1.json > gpurun_out/r2c_pin_sub_off0.5_il1.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 4 --methods pm,p3m,snapvspk --lattice-offset 0.25 --estimator-interlace 0 --out gpurun_out/r2c_pin_sub_off0.25_il0.json > gpurun_out/r2c_pin_sub_off0.25_il0.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 4 --methods p3m --lattice-offset 0.5 --estimator-interlace 0 --eta 0.0125 --max-sub 64 --out gpurun_out/r2c_pin_sub_eta0125.json > gpurun_out/r2c_pin_sub_eta0125.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 4 --methods p3m --lattice-offset 0.5 --estimator-interlace 0 --eta 0.05 --out gpurun_out/r2c_pin_sub_eta05.json > gpurun_out/r2c_pin_sub_eta05.log 2>&1
tail -n 12 gpurun_out/r2c_pin_*.log
