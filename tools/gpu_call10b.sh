#!/bin/bash
# 2-GPU call: dress rehearsal of the 8-GPU run (345M values files at real size, fp16 build), transposed MX quantiser,
# MoE N=2 after it, Mask R-CNN N=2 against the twin
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 300 python tools/kernel_check.py mx > gpurun_out/c10b_mx.log 2>&1; echo "mx rc=$?"; grep -h 'quantize_t\|"ok": false\|summary' gpurun_out/c10b_mx.log | cut -c1-300
EX=build/ref_examples/megatron-deepspeed/gpt2_345m
ROOT=/tmp/b2f_w
B2F_TRAIN_ITERS=30 B2F_OVERRIDE_ARGS="--log-interval 10 --eval-interval 100000 --save-interval 100000" timeout 400 \
  python -m b200flow.launcher install w2 pytorchjob-distributed -f $EX/pretrain-ddp-zero1.yaml --root $ROOT --max-restarts 0 \
  --set resources.nnodes=1 --set resources.nproc_per_node=2 > gpurun_out/c10b_w2.status 2> gpurun_out/c10b_w2.err
echo "w2 rc=$?"; tail -n 1 gpurun_out/c10b_w2.status | cut -c1-300
cat $ROOT/.b200flow/releases/w2/pods/*/stdout.log > gpurun_out/c10b_w2.log 2>/dev/null
grep -h "iteration .*tokens per second" gpurun_out/c10b_w2.log | tail -n 3 | cut -c1-420; tail -n 5 gpurun_out/c10b_w2.log | cut -c1-300
B2F_TRAIN_ITERS=30 B2F_OVERRIDE_ARGS="--log-interval 10 --eval-interval 100000 --save-interval 100000 --pipeline-model-parallel-size 1" timeout 400 \
  python -m b200flow.launcher install w3 pytorchjob-distributed -f $EX/pretrain-ddp-tp-pp-zero1.yaml --root $ROOT --max-restarts 0 \
  --set resources.nnodes=1 --set resources.nproc_per_node=2 > gpurun_out/c10b_w3.status 2> gpurun_out/c10b_w3.err
echo "w3(tp2) rc=$?"; tail -n 1 gpurun_out/c10b_w3.status | cut -c1-300
cat $ROOT/.b200flow/releases/w3/pods/*/stdout.log > gpurun_out/c10b_w3.log 2>/dev/null
grep -h "iteration .*tokens per second" gpurun_out/c10b_w3.log | tail -n 3 | cut -c1-420; tail -n 5 gpurun_out/c10b_w3.log | cut -c1-300
B2F_HALF=fp16 timeout 300 $P --master-port 29751 tools/multigpu_check.py fused_gemm tp > gpurun_out/c10b_fp16_tp.log 2>&1; echo "fp16 fused tp rc=$?"
grep -h '"ok": false\|summary' gpurun_out/c10b_fp16_tp.log | cut -c1-300 | head
timeout 300 $P --master-port 29753 bench.py --gpus 2 --workload gpt-moe-1.3b --steps 8 --warmup 6 --baseline none > gpurun_out/c10b_moe2.json 2> gpurun_out/c10b_moe2.err; echo "moe2 rc=$?"
tail -n 1 gpurun_out/c10b_moe2.json | cut -c1-700
timeout 300 $P --master-port 29755 bench.py --gpus 2 --workload maskrcnn --steps 8 --warmup 4 --baseline-steps 8 > gpurun_out/c10b_mrcnn2.json 2> gpurun_out/c10b_mrcnn2.err; echo "mrcnn2 rc=$?"
tail -n 1 gpurun_out/c10b_mrcnn2.json | cut -c1-1400
