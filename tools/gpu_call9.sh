#!/bin/bash
# 1-GPU call: Mask R-CNN (fixed-shape path) vs the torchvision twin; ncu --set full of the round-2 kernels
mkdir -p gpurun_out
timeout 300 python bench.py --workload maskrcnn --steps 8 --warmup 4 --baseline none > gpurun_out/c9_mrcnn.json 2> gpurun_out/c9_mrcnn.err; echo "mrcnn rc=$?"
B2F_MASKRCNN_STATIC=0 timeout 300 python bench.py --workload maskrcnn --steps 8 --warmup 4 --baseline none > gpurun_out/c9_mrcnn_dyn.json 2> gpurun_out/c9_mrcnn_dyn.err; echo "mrcnn dynamic rc=$?"
timeout 300 python bench.py --workload maskrcnn --impl twin --steps 8 --warmup 4 > gpurun_out/c9_mrcnn_twin.json 2> gpurun_out/c9_mrcnn_twin.err; echo "mrcnn twin rc=$?"
for f in gpurun_out/c9_mrcnn.json gpurun_out/c9_mrcnn_dyn.json gpurun_out/c9_mrcnn_twin.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"],1), d["unit"], "ms/step", round(d["ms_per_step"],1), d["clocks"].get("power_w_max"), d.get("loss_first_last"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
tail -n 3 gpurun_out/c9_mrcnn.err | cut -c1-300
python tools/prof_r2.py > gpurun_out/c9_prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'b2f_gemm_mx_kernel|b2f_moe_up_kernel|b2f_moe_down_kernel|b2f_ag_gemm2_kernel|b2f_gemm_rs2_kernel|attn_fwd2_kernel' -c 14 -o gpurun_out/c9_r2 python tools/prof_r2.py > gpurun_out/c9_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/c9_ncu.log; ls -la gpurun_out/c9_r2.ncu-rep
