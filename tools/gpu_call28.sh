#!/bin/bash
# 1-GPU call: the driver's smoke() on the final libraries, block-scaled GEMM checks after the last rebuild
mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c28_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/c28_smoke.log | cut -c1-400
timeout 200 python tools/kernel_check.py mx k5 > gpurun_out/c28_mx_k5.log 2>&1; echo "mx k5 rc=$?"; grep -h '"ok": false\|summary' gpurun_out/c28_mx_k5.log | cut -c1-300
