#!/bin/bash
# 2-GPU call: the cross-GPU protocols (K1 fp32 / bf16 wire, K2 peer / multimem, K3, K4, K4', K5, MoE EP, TP engine) under the race-delay build
mkdir -p gpurun_out
bash tools/race_check.sh 2 1 > gpurun_out/c27_race.log 2>&1; echo "race rc=$?"; cat gpurun_out/c27_race.log | cut -c1-300
