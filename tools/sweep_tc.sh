#!/bin/bash
# Rebuilds the tensor-core kernel with different compile-time layouts on the GPU box and times each (tuning aid).
cd "$(dirname "$0")/.."
PKG=distributed-inference-system-based-storm_b200
for cfg in "$@"; do
  set -- $cfg
  export SIB_NVCC_EXTRA="-DSIB_CONV1_WARPS=$1 -DSIB_BATCH_ROWS=$2 -DSIB_SPLIT_ACC=$3 -DSIB_CONV2_WARPS=${4:-3}"
  touch $PKG/csrc/sib_kernels_tc.cu
  python $PKG/build.py > /dev/null 2>&1 || { echo "build failed $cfg"; continue; }
  python -m pytest tests/test_gpu_tc.py -m gpu -x -q --timeout 120 -k "matches_oracle or full_size" 2>&1 | tail -1
  for b in 65536; do
    python bench.py --steps 20 --warmup 3 --skip-e2e --skip-cpu --skip-latency --no-guard --batch $b 2>/dev/null | python -c "import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('cfg $cfg batch $b', round(d['value']/1e6,2), 'Mimg/s frac', round(d['roofline']['frac'],4))"
  done
done
