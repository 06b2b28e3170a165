#!/bin/bash
# 2-GPU call: Mask R-CNN DP2 on the final code (the N=2 point of config 3)
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 300 $P --master-port 29851 bench.py --gpus 2 --workload maskrcnn --steps 12 --warmup 6 --baseline none > gpurun_out/c29_mrcnn2.json 2> gpurun_out/c29_mrcnn2.err; echo "mrcnn2 rc=$?"
tail -n 1 gpurun_out/c29_mrcnn2.json | cut -c1-900
