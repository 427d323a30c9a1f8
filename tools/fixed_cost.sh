#!/bin/bash
# Fixed cost of one launch of the fused kernel: kernel_ms over launch sizes of 1, 148, 296 ... pairs (16 images each).
# usage (GPU box): [SIZES="1 16 65536"] [PRESETS=cifar-lenet5] bash tools/fixed_cost.sh > gpurun_out/fixed_cost.jsonl
export SIB_BENCH_NO_SUSTAINED=1
SIZES=${SIZES:-"1 16 2368 4736 9472 18944 37888 65536 131072 262144"}
for preset in ${PRESETS:-cifar-lenet5 mnist-lenet5}; do
  for b in $SIZES; do
    python bench.py --preset $preset --batch $b --steps 50 --warmup 5 --skip-e2e --skip-cpu --skip-latency 2>/dev/null |
      python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print(json.dumps({'preset':'$preset','images_per_launch':$b,'kernel_us':round(r['kernel_ms']*1e3,2),'guard_us':round(r['guard_kernel_ms']*1e3,2),'step_us':round(r['step_ms']*1e3,2),'frac':round(r['frac'],4)}))"
  done
done
