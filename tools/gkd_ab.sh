#!/bin/bash
# Developer A/B of the fused gather+kick+drift tail's launch bound (CTAs per SM): rebuild gather.o on the box, read the
# gather phase and the step time from bench.py.
cd "$(dirname "$0")/../nbody_b200/csrc" || exit 1
for c in "$@"; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC \
      -DGKD_MIN_CTAS=$c -c gather.cu -o gather.o 2>/dev/null || exit 1
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libnbody_b200.so *.o \
      -L/usr/local/cuda/lib64 -lcufft -Xlinker -rpath -Xlinker /usr/local/cuda/lib64 || exit 1
  (cd ../.. && timeout 200 python bench.py --no-cpu --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ctas=$c', 'ms_per_step', round(d['ms_per_step'],2), 'gather', d['phase_ms']['gather'])")
done
