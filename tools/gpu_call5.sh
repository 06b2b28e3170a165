#!/bin/bash
# 2-GPU call: new K3 (gather clusters rejoin) / K4 (interleaved order) — protocol checks, perf at the Llama shapes, TP2 / PP2 bench points
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 200 python tools/kernel_check.py attn_drop attn_bwd > gpurun_out/c5_attn.jsonl 2>&1; echo "attn_drop rc=$?"
timeout 300 $P --master-port 29741 tools/multigpu_check.py fused_gemm tp_replicated tp > gpurun_out/c5_fused.log 2>&1; echo "fused_gemm rc=$?"
timeout 300 $P --master-port 29743 tools/tp_perf.py 16 24 32 > gpurun_out/c5_tp_perf_new.jsonl 2> gpurun_out/c5_tp_perf_new.err; echo "tp_perf new rc=$?"
B2F_RS_INTERLEAVE=0 B2F_AG_REJOIN=0 timeout 200 $P --master-port 29745 tools/tp_perf.py 16 > gpurun_out/c5_tp_perf_old.jsonl 2> gpurun_out/c5_tp_perf_old.err; echo "tp_perf old rc=$?"
timeout 400 $P --master-port 29749 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/c5_tp2.json 2> gpurun_out/c5_tp2.err; echo "tp2 rc=$?"
timeout 200 $P --master-port 29751 bench.py --gpus 2 --tp 1 --pp 2 --steps 3 --warmup 3 --baseline none > gpurun_out/c5_pp2_zb.json 2> gpurun_out/c5_pp2_zb.err; echo "pp2 zb rc=$?"
timeout 200 $P --master-port 29753 bench.py --gpus 2 --tp 1 --pp 2 --steps 3 --warmup 3 --baseline none --no-zero-bubble > gpurun_out/c5_pp2_1f1b.json 2> gpurun_out/c5_pp2_1f1b.err; echo "pp2 1f1b rc=$?"
timeout 300 $P --master-port 29755 bench.py --gpus 2 --workload gpt-moe-1.3b --steps 6 --warmup 10 --baseline none > gpurun_out/c5_moe_k5.json 2> gpurun_out/c5_moe_k5.err; echo "moe k5 rc=$?"
grep -h '"ok": false' gpurun_out/c5_attn.jsonl gpurun_out/c5_fused.log | cut -c1-420
grep -h 'cost\|summary' gpurun_out/c5_attn.jsonl gpurun_out/c5_fused.log | cut -c1-300
python - <<'PY'
import json
def load(f):
    out={}
    try:
        for l in open(f):
            if l.startswith("{"):
                d=json.loads(l); out[(d["shape"], d["comm_ctas"])]=d
    except Exception as e: print(f, e)
    return out
new, old = load("gpurun_out/c5_tp_perf_new.jsonl"), load("gpurun_out/c5_tp_perf_old.jsonl")
for (shape, c), d in sorted(new.items()):
    o = old.get((shape, 16), {})
    print(f"{shape:20s} ctas {c:2d} fused {d['fused_ms']:.4f} gemm {d['gemm_alone_ms']:.4f} ratio {d['fused_over_gemm']:.3f} | old fused {o.get('fused_ms')} ratio {o.get('fused_over_gemm')} nccl {d['nccl_plus_gemm_ms']:.4f}")
PY
for f in gpurun_out/c5_tp2.json gpurun_out/c5_pp2_zb.json gpurun_out/c5_pp2_1f1b.json gpurun_out/c5_moe_k5.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"]), d["config"].get("parallelism"), "vs_baseline", d.get("vs_baseline"), d.get("loss_parity"), (d.get("baseline") or {}).get("value"), d.get("loss_curve"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
for f in gpurun_out/c5_*.err; do echo "== $f"; tail -n 3 "$f" | cut -c1-300; done
