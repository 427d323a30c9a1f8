#!/bin/bash
# Times the guarded step for each "threads images" pair given (guard kernel compile-time shape).
cd "$(dirname "$0")/.."
PKG=distributed-inference-system-based-storm_b200
for cfg in "$@"; do
  set -- $cfg
  export SIB_NVCC_EXTRA="-DSIB_GUARD_THREADS=$1 -DSIB_GUARD_IMGS=$2"
  touch $PKG/csrc/sib_kernels_guard.cu
  python $PKG/build.py > /tmp/build.log 2>&1 || { echo "build failed: $cfg"; tail -5 /tmp/build.log; continue; }
  for b in 65536 16384; do
  timeout 120 python bench.py --steps 20 --warmup 3 --skip-e2e --skip-cpu --skip-latency --batch $b 2>/dev/null | python -c "import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('[$cfg] batch $b', round(d['value']/1e6,2), 'Mimg/s  guard ms', round(d['roofline'].get('guard_kernel_ms',0),4))"
  done
done
