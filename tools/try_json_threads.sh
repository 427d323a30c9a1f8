#!/bin/bash
# Decode-kernel CTA size: single-tuple round trip and saturated tuple rate for each thread count given.
cd "$(dirname "$0")/.."
PKG=distributed-inference-system-based-storm_b200
for t in "$@"; do
  export SIB_NVCC_EXTRA="-DSIB_JSON_THREADS=$t"
  touch $PKG/csrc/sib_kernels_json.cu
  python $PKG/build.py > /tmp/build.log 2>&1 || { echo "build failed: $t"; tail -5 /tmp/build.log; continue; }
  echo "[$t] $(timeout 200 python -m pytest tests/test_gpu_json.py -m gpu -q --timeout 120 2>&1 | tail -1)"
  echo "[$t] $(timeout 100 python tools/single_shot.py | grep 'gpu decode')"
  echo "[$t] $(timeout 200 python tools/tuple_rate.py | grep 'gpu decode' | tail -2 | tr '\n' ' ')"
done
