#!/bin/bash
# Runs prebuilt variants of libsib.so (distributed-inference-system-based-storm_b200/variants/libsib_*.so, built here with
# -D flags) on the GPU box: a parity subset and the kernel time at 65 536 images each; restores the default library.
# usage (GPU box): [FULL=1] bash tools/try_variants.sh > gpurun_out/variants.log 2>&1
P=distributed-inference-system-based-storm_b200
cp $P/libsib.so /tmp/libsib_default.so
for v in $P/variants/libsib_*.so; do
  echo "=== $v"
  cp $v $P/libsib.so
  if [ -n "$FULL" ]; then python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -1
  else python -m pytest tests/test_gpu_tc.py -x -q -m gpu -k "launch_size or batch_edges or matches_oracle or guard_recomputes" 2>&1 | tail -1; fi
  SIZES="${SIZES:-65536}" PRESETS="${PRESETS:-cifar-lenet5}" bash tools/fixed_cost.sh
done
cp /tmp/libsib_default.so $P/libsib.so
