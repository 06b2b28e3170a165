#!/bin/bash
# 8-GPU call C: pretrain-ddp-tp-pp-zero1.yaml (TP2 x PP2 x DP2, fp16, one micro-batch per step) after the pipeline
# boundary connect fix
mkdir -p gpurun_out
EX=build/ref_examples/megatron-deepspeed/gpt2_345m
ROOT=/tmp/b2f_w
W=pretrain-ddp-tp-pp-zero1
B2F_TRAIN_ITERS=40 B2F_OVERRIDE_ARGS="--log-interval 10 --eval-interval 100000 --save-interval 100000" timeout 420 \
  python -m b200flow.launcher install w-$W pytorchjob-distributed -f $EX/$W.yaml --root $ROOT --max-restarts 0 \
  --set resources.nnodes=1 --set resources.nproc_per_node=8 --set 'resources.limits.nvidia\.com/gpu=8' \
  --set 'resources.requests.nvidia\.com/gpu=8' > gpurun_out/c20_$W.status 2> gpurun_out/c20_$W.err
echo "$W rc=$?"; tail -n 1 gpurun_out/c20_$W.status | cut -c1-300
cat $ROOT/.b200flow/releases/w-$W/pods/*/stdout.log > gpurun_out/c20_$W.log 2>/dev/null
grep -h "iteration .*tokens per second" gpurun_out/c20_$W.log | tail -n 4 | cut -c1-420
grep -h "Error\|error\|WARN" gpurun_out/c20_$W.log | head -n 4 | cut -c1-300
