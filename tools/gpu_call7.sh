#!/bin/bash
mkdir -p gpurun_out
P="python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1"
timeout 120 python tools/k4_loopback_perf.py > gpurun_out/c7_loop.jsonl 2>&1; echo "loopback rc=$?"
timeout 300 $P --master-port 29741 tools/multigpu_check.py fused_gemm tp > gpurun_out/c7_fused.log 2>&1; echo "fused_gemm rc=$?"
timeout 200 $P --master-port 29743 tools/tp_perf.py 16 > gpurun_out/c7_tp_perf.jsonl 2> gpurun_out/c7_tp_perf.err; echo "tp_perf rc=$?"
B2F_RS_INTERLEAVE=0 timeout 200 $P --master-port 29745 tools/tp_perf.py 16 > gpurun_out/c7_tp_perf_nointer.jsonl 2> gpurun_out/c7_tp_perf_nointer.err; echo "tp_perf nointer rc=$?"
timeout 300 $P --master-port 29747 tools/multigpu_check.py arena allreduce zero1 > gpurun_out/c7_allreduce.log 2>&1; echo "allreduce rc=$?"
timeout 300 $P --master-port 29749 tools/multigpu_check.py zero1_perf > gpurun_out/c7_zero1_perf.log 2>&1; echo "zero1_perf rc=$?"
timeout 400 $P --master-port 29751 bench.py --gpus 2 --steps 4 --warmup 3 --baseline none > gpurun_out/c7_tp2.json 2> gpurun_out/c7_tp2.err; echo "tp2 rc=$?"
timeout 300 $P --master-port 29753 bench.py --gpus 2 --workload bert-large --steps 6 --warmup 3 > gpurun_out/c7_bert2.json 2> gpurun_out/c7_bert2.err; echo "bert2 rc=$?"
cat gpurun_out/c7_loop.jsonl | cut -c1-220
grep -h '"ok": false\|summary' gpurun_out/c7_fused.log gpurun_out/c7_allreduce.log gpurun_out/c7_zero1_perf.log | cut -c1-400
grep -h 'perf\|nccl_bf16' gpurun_out/c7_allreduce.log gpurun_out/c7_zero1_perf.log | grep '"rank": 0' | cut -c1-330
python - <<'PY'
import json
def load(f):
    out={}
    try:
        for l in open(f):
            if l.startswith("{"):
                d=json.loads(l); out[d["shape"]]=d
    except Exception as e: print(f, e)
    return out
a, b = load("gpurun_out/c7_tp_perf.jsonl"), load("gpurun_out/c7_tp_perf_nointer.jsonl")
for k, d in a.items():
    o = b.get(k, {})
    print(f"{k:20s} fused {d['fused_ms']:.4f} gemm {d['gemm_alone_ms']:.4f} ratio {d['fused_over_gemm']:.3f} | no-interleave fused {o.get('fused_ms')} | nccl+gemm {d['nccl_plus_gemm_ms']:.4f}")
PY
for f in gpurun_out/c7_tp2.json gpurun_out/c7_bert2.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"]), d["config"].get("parallelism"), "vs_baseline", d.get("vs_baseline"), (d.get("baseline") or {}).get("value"), d.get("loss_curve") or d.get("loss_first_last"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
for f in gpurun_out/c7_*.err; do echo "== $f"; tail -n 2 "$f" | cut -c1-300; done
