#!/bin/bash
# 1-GPU call: Mask R-CNN after the loss fix (fixed-shape / variable-length / torchvision twin), kernel shares of the
# Mask R-CNN step and of the fp8 MoE layer
mkdir -p gpurun_out
timeout 300 python bench.py --workload maskrcnn --steps 8 --warmup 4 --baseline none > gpurun_out/c10_mrcnn.json 2> gpurun_out/c10_mrcnn.err; echo "mrcnn rc=$?"
B2F_MASKRCNN_STATIC=0 timeout 300 python bench.py --workload maskrcnn --steps 8 --warmup 4 --baseline none > gpurun_out/c10_mrcnn_dyn.json 2> gpurun_out/c10_mrcnn_dyn.err; echo "mrcnn dynamic rc=$?"
timeout 300 python bench.py --workload maskrcnn --impl twin --steps 8 --warmup 4 > gpurun_out/c10_mrcnn_twin.json 2> gpurun_out/c10_mrcnn_twin.err; echo "mrcnn twin rc=$?"
for f in gpurun_out/c10_mrcnn.json gpurun_out/c10_mrcnn_dyn.json gpurun_out/c10_mrcnn_twin.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"],1), d["unit"], "ms/step", round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"],1), d.get("loss_first_last"), d.get("loss_after_warmup"), d.get("invalid"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
for w in maskrcnn maskrcnn_dyn moe_layer; do timeout 300 python tools/prof_torch.py $w > gpurun_out/c10_shares_$w.txt 2> gpurun_out/c10_shares_$w.err; echo "shares $w rc=$?"; head -24 gpurun_out/c10_shares_$w.txt | cut -c1-150; done
