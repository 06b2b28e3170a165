#!/bin/bash
# 4-GPU call: the N=4 points of configs 2, 3 and 5 (own arm; the stock arms were timed at 1 / 2 / 8)
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=4 --master-addr 127.0.0.1"
timeout 300 $P --master-port 29841 bench.py --gpus 4 --workload bert-large --steps 12 --warmup 5 --baseline none > gpurun_out/c22_bert4.json 2> gpurun_out/c22_bert4.err; echo "bert4 rc=$?"
tail -n 1 gpurun_out/c22_bert4.json | cut -c1-600
timeout 300 $P --master-port 29843 bench.py --gpus 4 --workload maskrcnn --steps 12 --warmup 6 --baseline none > gpurun_out/c22_mrcnn4.json 2> gpurun_out/c22_mrcnn4.err; echo "mrcnn4 rc=$?"
tail -n 1 gpurun_out/c22_mrcnn4.json | cut -c1-600
timeout 300 $P --master-port 29845 bench.py --gpus 4 --workload gpt-moe-1.3b --steps 8 --warmup 8 --baseline none > gpurun_out/c22_moe4.json 2> gpurun_out/c22_moe4.err; echo "moe4 rc=$?"
tail -n 1 gpurun_out/c22_moe4.json | cut -c1-600
