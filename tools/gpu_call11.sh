#!/bin/bash
# 8-GPU call A: GPT-MoE 1.3B (K5, EP8) against the stock twin on the same box; the two GPT-2 345M values files of the
# reference through the launcher at their real size; K1 / K2 against NCCL at 8 GPUs
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1"
timeout 420 $P --master-port 29811 bench.py --gpus 8 --workload gpt-moe-1.3b --steps 10 --warmup 6 --baseline-steps 10 > gpurun_out/c11_moe8.json 2> gpurun_out/c11_moe8.err; echo "moe8 rc=$?"
tail -n 1 gpurun_out/c11_moe8.json | cut -c1-1500
EX=build/ref_examples/megatron-deepspeed/gpt2_345m
ROOT=/tmp/b2f_w
for W in pretrain-ddp-zero1 pretrain-ddp-tp-pp-zero1; do
  B2F_TRAIN_ITERS=40 B2F_OVERRIDE_ARGS="--log-interval 10 --eval-interval 100000 --save-interval 100000" timeout 420 \
    python -m b200flow.launcher install w-$W pytorchjob-distributed -f $EX/$W.yaml --root $ROOT --max-restarts 0 \
    --set resources.nnodes=1 --set resources.nproc_per_node=8 > gpurun_out/c11_$W.status 2> gpurun_out/c11_$W.err
  echo "$W rc=$?"; tail -n 1 gpurun_out/c11_$W.status | cut -c1-300
  cat $ROOT/.b200flow/releases/w-$W/pods/*/stdout.log > gpurun_out/c11_$W.log 2>/dev/null
  grep -h "iteration .*tokens per second\|native\|half type\|loss scale" gpurun_out/c11_$W.log | tail -n 6 | cut -c1-420
done
timeout 300 $P --master-port 29813 tools/multigpu_check.py arena allreduce zero1 > gpurun_out/c11_allreduce.log 2>&1; echo "allreduce rc=$?"
timeout 300 $P --master-port 29815 tools/multigpu_check.py zero1_perf > gpurun_out/c11_zero1_perf.log 2>&1; echo "zero1_perf rc=$?"
grep -h '"ok": false\|summary' gpurun_out/c11_allreduce.log gpurun_out/c11_zero1_perf.log | cut -c1-400
grep -h 'perf\|nccl_bf16' gpurun_out/c11_allreduce.log gpurun_out/c11_zero1_perf.log | grep '"rank": 0' | cut -c1-300
