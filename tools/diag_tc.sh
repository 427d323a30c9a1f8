#!/bin/bash
# Rebuilds the tensor-core kernel with diagnostic / tuning macros on the GPU box and times each build (guard off).
# usage: tools/diag_tc.sh "<nvcc -D flags>" ...      ("" = the default build)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; exec > >(tee -a gpurun_out/diag.log) 2>&1
PKG=distributed-inference-system-based-storm_b200
for cfg in "$@"; do
  export SIB_NVCC_EXTRA="$cfg"
  touch $PKG/csrc/sib_kernels_tc.cu
  python $PKG/build.py > /tmp/build.log 2>&1 || { echo "build failed: $cfg"; tail -5 /tmp/build.log; continue; }
  for b in ${BATCHES:-65536}; do
    SIB_BENCH_NOCHECK=1 timeout 120 python bench.py --steps 20 --warmup 3 --skip-e2e --skip-cpu --skip-latency --no-guard --batch $b 2>/dev/null | tail -1 | python -c "import sys,json
d=json.loads(sys.stdin.read()); print('cfg [$cfg] batch $b:', round(d['value']/1e6,2), 'Mimg/s  kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4))"
  done
done
export SIB_NVCC_EXTRA=""
touch $PKG/csrc/sib_kernels_tc.cu
python $PKG/build.py > /dev/null 2>&1
