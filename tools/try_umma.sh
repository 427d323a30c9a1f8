#!/bin/bash
# Builds libsib.so with conv1 on tcgen05 (SIB_CONV1_UMMA=1) on the GPU box, runs the tensor-core tests and times it.
cd "$(dirname "$0")/.."
PKG=distributed-inference-system-based-storm_b200
export SIB_NVCC_EXTRA="-DSIB_CONV1_UMMA=1 $1"
touch $PKG/csrc/sib_kernels_tc.cu
python $PKG/build.py > /tmp/build.log 2>&1 || { tail -20 /tmp/build.log; exit 1; }
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q --timeout 60 2>&1 | tail -15
for b in 16384 65536; do
  timeout 120 python bench.py --steps 20 --warmup 3 --skip-e2e --skip-cpu --skip-latency --no-guard --batch $b 2>/dev/null | python -c "import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('umma batch $b', round(d['value']/1e6,2), 'Mimg/s frac', round(d['roofline']['frac'],4))"
done
