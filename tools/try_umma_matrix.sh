#!/bin/bash
# A/B matrix over the tcgen05 conv1 build switches (timing only; parity is tools/try_umma.sh's job).
cd "$(dirname "$0")/.."
PKG=distributed-inference-system-based-storm_b200
for flags in "$@"; do
  export SIB_NVCC_EXTRA="-DSIB_CONV1_UMMA=1 $flags"
  touch $PKG/csrc/sib_kernels_tc.cu
  python $PKG/build.py > /tmp/build.log 2>&1 || { echo "build failed: $flags"; tail -5 /tmp/build.log; continue; }
  timeout 120 python bench.py --steps 20 --warmup 3 --skip-e2e --skip-cpu --skip-latency --no-guard --batch 65536 2>/dev/null | python -c "import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('[$flags]', round(d['value']/1e6,2), 'Mimg/s')"
done
