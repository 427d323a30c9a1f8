#!/bin/bash
# Runs both tcgen05 probes on the GPU box, one risky case per process, and collects their JSON lines.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
( timeout 120 tools/tcgen05_probe ) > gpurun_out/tcgen05_probe_r01.jsonl 2>&1
out=gpurun_out/tcgen05_probe2_r01.jsonl; : > $out
for d in 0 8 6 12 22 2; do timeout 60 tools/tcgen05_probe2 ss $d >> $out 2>&1 || echo "{\"probe\": \"ss $d\", \"exit\": $?}" >> $out; done
for c in 6 3; do timeout 60 tools/tcgen05_probe2 ldst $c >> $out 2>&1 || echo "{\"probe\": \"ldst $c\", \"exit\": $?}" >> $out; done
for s in 0 6; do timeout 60 tools/tcgen05_probe2 rate $s >> $out 2>&1 || echo "{\"probe\": \"rate $s\", \"exit\": $?}" >> $out; done
cat gpurun_out/tcgen05_probe_r01.jsonl $out
