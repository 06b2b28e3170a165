#!/bin/bash
# Run kernel bring-up groups, each in its own process under a timeout.
# usage: tools/run_checks.sh group1 group2 ...
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for g in "$@"; do
  echo "=== $g"
  timeout 240 python tools/kernel_check.py $g > gpurun_out/check_$g.log 2>&1
  echo "exit=$? $(tail -n 1 gpurun_out/check_$g.log | cut -c1-300)"
done
