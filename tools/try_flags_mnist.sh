#!/bin/bash
# Like try_flags_tested.sh, timing mnist-lenet5 as well as cifar-lenet5 (65536 images per launch, guard off).
cd "$(dirname "$0")/.."
PKG=distributed-inference-system-based-storm_b200
for flags in "$@"; do
  export SIB_NVCC_EXTRA="$flags"
  touch $PKG/csrc/sib_kernels_tc.cu
  python $PKG/build.py > /tmp/build.log 2>&1 || { echo "build failed: $flags"; tail -5 /tmp/build.log; continue; }
  echo "[$flags] $(timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q --timeout 60 2>&1 | tail -1)"
  for p in mnist-lenet5 cifar-lenet5; do
  timeout 120 python bench.py --steps 20 --warmup 3 --skip-e2e --skip-cpu --skip-latency --no-guard --preset $p 2>/dev/null | python -c "import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('[$flags] $p', round(d['value']/1e6,2), 'Mimg/s')"
  done
done
