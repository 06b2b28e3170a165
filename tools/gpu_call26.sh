#!/bin/bash
# 1-GPU call: ncu --set full of the final attention kernels (forward with P published in quarters; backward; dropout forms)
mkdir -p gpurun_out
python tools/prof_r2.py attn > gpurun_out/c26_prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'attn_' -c 14 -o gpurun_out/c26_attn python tools/prof_r2.py attn > gpurun_out/c26_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/c26_ncu.log; ls -la gpurun_out/c26_attn.ncu-rep
