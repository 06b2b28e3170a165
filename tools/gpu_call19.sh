#!/bin/bash
# 8-GPU call B: the two GPT-2 345M values files of the reference through the launcher at their real size (GPU limit raised
# to 8 per pod); GPT-MoE EP8 on Zipf tokens with the twin; Mask R-CNN and BERT-large DP8 with their twins
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1"
EX=build/ref_examples/megatron-deepspeed/gpt2_345m
ROOT=/tmp/b2f_w
for W in pretrain-ddp-zero1 pretrain-ddp-tp-pp-zero1; do
  B2F_TRAIN_ITERS=40 B2F_OVERRIDE_ARGS="--log-interval 10 --eval-interval 100000 --save-interval 100000" timeout 420 \
    python -m b200flow.launcher install w-$W pytorchjob-distributed -f $EX/$W.yaml --root $ROOT --max-restarts 0 \
    --set resources.nnodes=1 --set resources.nproc_per_node=8 --set 'resources.limits.nvidia\.com/gpu=8' \
    --set 'resources.requests.nvidia\.com/gpu=8' > gpurun_out/c19_$W.status 2> gpurun_out/c19_$W.err
  echo "$W rc=$?"; tail -n 1 gpurun_out/c19_$W.status | cut -c1-300
  cat $ROOT/.b200flow/releases/w-$W/pods/*/stdout.log > gpurun_out/c19_$W.log 2>/dev/null
  grep -h "iteration .*tokens per second" gpurun_out/c19_$W.log | tail -n 3 | cut -c1-420
  grep -h "Error\|error" gpurun_out/c19_$W.log | head -n 3 | cut -c1-300
done
timeout 420 $P --master-port 29821 bench.py --gpus 8 --workload gpt-moe-1.3b --steps 8 --warmup 10 --baseline-steps 8 > gpurun_out/c19_moe8.json 2> gpurun_out/c19_moe8.err; echo "moe8 rc=$?"
tail -n 1 gpurun_out/c19_moe8.json | cut -c1-2400
timeout 420 $P --master-port 29823 bench.py --gpus 8 --workload maskrcnn --steps 10 --warmup 5 --baseline-steps 8 > gpurun_out/c19_mrcnn8.json 2> gpurun_out/c19_mrcnn8.err; echo "mrcnn8 rc=$?"
tail -n 1 gpurun_out/c19_mrcnn8.json | cut -c1-1800
timeout 420 $P --master-port 29825 bench.py --gpus 8 --workload bert-large --steps 10 --warmup 5 --baseline-steps 8 > gpurun_out/c19_bert8.json 2> gpurun_out/c19_bert8.err; echo "bert8 rc=$?"
tail -n 1 gpurun_out/c19_bert8.json | cut -c1-1800
