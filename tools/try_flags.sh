#!/bin/bash
# Times the tensor-core kernel (guard off, 65536 and 16384 images per launch) for each set of extra nvcc flags given.
cd "$(dirname "$0")/.."
PKG=distributed-inference-system-based-storm_b200
for flags in "$@"; do
  export SIB_NVCC_EXTRA="$flags"
  touch $PKG/csrc/sib_kernels_tc.cu
  python $PKG/build.py > /tmp/build.log 2>&1 || { echo "build failed: $flags"; tail -5 /tmp/build.log; continue; }
  for b in 65536 16384; do
  timeout 120 python bench.py --steps 20 --warmup 3 --skip-e2e --skip-cpu --skip-latency --no-guard --batch $b 2>/dev/null | python -c "import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('[$flags] batch $b', round(d['value']/1e6,2), 'Mimg/s')"
  done
done
