#!/bin/bash
# 4-GPU call: the bench's TP2 x PP2 point (fused TP kernels + split-backward pipeline schedule, 16 micro-batches) on the final code
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=4 --master-addr 127.0.0.1"
timeout 150 $P --master-port 29861 bench.py --gpus 4 --steps 2 --warmup 3 --baseline none > gpurun_out/c30_llama4.json 2> gpurun_out/c30_llama4.err; echo "llama4 rc=$?"
tail -n 1 gpurun_out/c30_llama4.json | cut -c1-1500; tail -n 3 gpurun_out/c30_llama4.err | cut -c1-300
