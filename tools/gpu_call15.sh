#!/bin/bash
# 2-GPU call: benches on Zipf-distributed tokens (loss curves that mean something): GPT-MoE EP2 with the twin, Llama TP2 with the NCCL arm
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 420 $P --master-port 29911 bench.py --gpus 2 --workload gpt-moe-1.3b --steps 10 --warmup 8 --baseline-steps 10 > gpurun_out/c15_moe2.json 2> gpurun_out/c15_moe2.err; echo "moe2 rc=$?"
tail -n 1 gpurun_out/c15_moe2.json | cut -c1-2200
timeout 500 $P --master-port 29913 bench.py --gpus 2 --steps 4 --warmup 3 --baseline-steps 4 > gpurun_out/c15_llama2.json 2> gpurun_out/c15_llama2.err; echo "llama2 rc=$?"
tail -n 1 gpurun_out/c15_llama2.json | cut -c1-2200
