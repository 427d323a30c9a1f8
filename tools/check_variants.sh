#!/bin/bash
# Builds the alternative tensor-core kernels (-DSIB_CONV2_UMMA=0: round-1 all-mma.sync; -DSIB_CONV1_UMMA=1: conv1 on tcgen05 through
# registers) on the GPU box, runs the tensor-core parity tests on each and times it; restores the default build at the end.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
PKG=distributed-inference-system-based-storm_b200
for cfg in "-DSIB_CONV2_UMMA=0" "-DSIB_CONV1_UMMA=1"; do
  export SIB_NVCC_EXTRA="$cfg"
  touch $PKG/csrc/sib_kernels_tc.cu
  python $PKG/build.py > /tmp/build.log 2>&1 || { echo "build failed: $cfg"; tail -5 /tmp/build.log; continue; }
  echo "== $cfg"
  timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q --timeout 100 -k "lenet5 or batch_edges or full_size or near_tie" 2>&1 | tail -2
  timeout 120 python bench.py --steps 20 --warmup 3 --skip-e2e --skip-cpu --skip-latency --no-guard 2>/dev/null | tail -1 | python -c "import sys,json
d=json.loads(sys.stdin.read()); print('  ', round(d['value']/1e6,1), 'M img/s, frac', round(d['roofline']['frac'],4))"
done
export SIB_NVCC_EXTRA=""
touch $PKG/csrc/sib_kernels_tc.cu
python $PKG/build.py > /dev/null 2>&1
