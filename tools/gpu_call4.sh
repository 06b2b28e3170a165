#!/bin/bash
# 2-GPU call: K5 across GPUs, K3/K4 at the Llama TP2 shapes (gather-CTA sweep), PP2 timelines, TP2 bench (fused vs NCCL), MoE loss curves
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 300 $P --master-port 29741 tools/multigpu_check.py moe_k5 > gpurun_out/c4_moe_k5.log 2>&1; echo "moe_k5 rc=$?"
timeout 300 $P --master-port 29743 tools/tp_perf.py 16 8 4 > gpurun_out/c4_tp_perf.jsonl 2> gpurun_out/c4_tp_perf.err; echo "tp_perf rc=$?"
B2F_PP_TRACE=gpurun_out/c4_trace_zb timeout 200 $P --master-port 29745 bench.py --gpus 2 --tp 1 --pp 2 --steps 1 --warmup 3 --baseline none > gpurun_out/c4_pp2_zb.json 2> gpurun_out/c4_pp2_zb.err; echo "pp2 zb rc=$?"
B2F_PP_TRACE=gpurun_out/c4_trace_1f1b timeout 200 $P --master-port 29747 bench.py --gpus 2 --tp 1 --pp 2 --steps 1 --warmup 3 --baseline none --no-zero-bubble > gpurun_out/c4_pp2_1f1b.json 2> gpurun_out/c4_pp2_1f1b.err; echo "pp2 1f1b rc=$?"
timeout 400 $P --master-port 29749 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/c4_tp2.json 2> gpurun_out/c4_tp2.err; echo "tp2 rc=$?"
timeout 300 $P --master-port 29751 bench.py --gpus 2 --workload gpt-moe-1.3b --steps 6 --warmup 8 --baseline none > gpurun_out/c4_moe_k5.json 2> gpurun_out/c4_moe_k5.err; echo "moe k5 rc=$?"
B2F_MOE_BF16=1 timeout 300 $P --master-port 29753 bench.py --gpus 2 --workload gpt-moe-1.3b --steps 6 --warmup 8 --baseline none > gpurun_out/c4_moe_bf16.json 2> gpurun_out/c4_moe_bf16.err; echo "moe bf16 rc=$?"
timeout 300 $P --master-port 29755 bench.py --gpus 2 --workload gpt-moe-1.3b --impl twin --steps 6 --warmup 8 > gpurun_out/c4_moe_twin.json 2> gpurun_out/c4_moe_twin.err; echo "moe twin rc=$?"
grep -h '"ok": false\|perf\|summary' gpurun_out/c4_moe_k5.log | cut -c1-500
cat gpurun_out/c4_tp_perf.jsonl | cut -c1-400
for f in gpurun_out/c4_pp2_zb.json gpurun_out/c4_pp2_1f1b.json gpurun_out/c4_tp2.json gpurun_out/c4_moe_k5.json gpurun_out/c4_moe_bf16.json gpurun_out/c4_moe_twin.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"]), d["config"].get("parallelism"), "vs_baseline", d.get("vs_baseline"), d.get("loss_parity"), (d.get("baseline") or {}).get("value"), d.get("loss_curve") or d.get("loss_first_last"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
tail -3 gpurun_out/c4_*.err | cut -c1-300 | tail -30
