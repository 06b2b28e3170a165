#!/bin/bash
# 1-GPU call: conv epilogue kernel, Mask R-CNN with the batched fixed-shape path (+ trunk graphs) against eager
mkdir -p gpurun_out
timeout 300 python tools/kernel_check.py detection maskrcnn_step > gpurun_out/c17_det.log 2>&1; echo "detection rc=$?"; grep -h '"ok": false\|summary' gpurun_out/c17_det.log | cut -c1-300
timeout 300 python bench.py --workload maskrcnn --steps 10 --warmup 5 --baseline none > gpurun_out/c17_mrcnn_graph.json 2> gpurun_out/c17_mrcnn_graph.err; echo "mrcnn graph rc=$?"
B2F_MASKRCNN_GRAPH=0 timeout 300 python bench.py --workload maskrcnn --steps 10 --warmup 5 --baseline none > gpurun_out/c17_mrcnn_eager.json 2> gpurun_out/c17_mrcnn_eager.err; echo "mrcnn eager rc=$?"
for f in gpurun_out/c17_mrcnn_graph.json gpurun_out/c17_mrcnn_eager.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"],1), d["unit"], "ms/step", round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"],1), d.get("loss_first_last"), d.get("loss_after_warmup"), d["config"].get("trunk_cuda_graph"), d["clocks"].get("power_w_max"), d.get("invalid"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
tail -n 4 gpurun_out/c17_mrcnn_graph.err | cut -c1-300
timeout 200 python tools/prof_torch.py maskrcnn 16 > gpurun_out/c17_shares_maskrcnn.txt 2>&1; head -19 gpurun_out/c17_shares_maskrcnn.txt | cut -c1-150
