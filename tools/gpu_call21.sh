#!/bin/bash
# 8-GPU call D: Mask R-CNN DP8 with the LR warm-up (the run in call 19 diverged at the full linearly-scaled rate)
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1"
timeout 300 $P --master-port 29831 bench.py --gpus 8 --workload maskrcnn --steps 12 --warmup 6 --baseline none > gpurun_out/c21_mrcnn8.json 2> gpurun_out/c21_mrcnn8.err; echo "mrcnn8 rc=$?"
tail -n 1 gpurun_out/c21_mrcnn8.json | cut -c1-1800
