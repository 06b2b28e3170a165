#!/bin/bash
# 2-GPU call: multi-GPU protocol tests, then TP2 and PP2 Llama-7B bench points (fused vs NCCL arm; zero-bubble vs 1F1B)
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 300 python tools/kernel_check.py attn_bwd fp8 gemm_tn gemm_tt > gpurun_out/c2_checks.jsonl 2>&1; echo "kernel checks rc=$?"
timeout 600 python -m pytest tests/test_multigpu.py -q 2>&1 | tail -15 > gpurun_out/c2_multigpu.log; tail -3 gpurun_out/c2_multigpu.log
timeout 200 $P --master-port 29741 tools/multigpu_check.py ulysses > gpurun_out/c2_ulysses.log 2>&1; echo "ulysses rc=$?"
timeout 500 $P --master-port 29742 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/c2_tp2.json 2> gpurun_out/c2_tp2.err; echo "tp2 rc=$?"
timeout 300 $P --master-port 29762 bench.py --gpus 2 --tp 1 --pp 2 --steps 4 --warmup 3 --baseline none > gpurun_out/c2_pp2_zb.json 2> gpurun_out/c2_pp2_zb.err; echo "pp2 zb rc=$?"
timeout 300 $P --master-port 29772 bench.py --gpus 2 --tp 1 --pp 2 --steps 4 --warmup 3 --baseline none --no-zero-bubble > gpurun_out/c2_pp2_1f1b.json 2> gpurun_out/c2_pp2_1f1b.err; echo "pp2 1f1b rc=$?"
grep -h '"ok": false' gpurun_out/c2_checks.jsonl | cut -c1-400
grep -h '"ok"' gpurun_out/c2_ulysses.log | cut -c1-300
for f in gpurun_out/c2_tp2.json gpurun_out/c2_pp2_zb.json gpurun_out/c2_pp2_1f1b.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["value"], d["config"]["parallelism"], "vs_baseline", d.get("vs_baseline"), d.get("loss_parity"), (d.get("baseline") or {}).get("value"), d["loss_first_last"])
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
