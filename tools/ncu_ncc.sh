#!/bin/bash
# ncu --set full of the patch-NCC kernels (ours and the compiled reference's) inside tools/ncc_bench.py.
set -u
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:"warp_patch_kernel|forward_mode" --launch-skip 8 -c 4 \
    -f -o gpurun_out/ncc python tools/ncc_bench.py --iters 3 > gpurun_out/ncu_ncc.log 2>&1
echo "ncu rc=$?"
python tools/ncu_summary.py gpurun_out/ncc.ncu-rep --lines 40 > gpurun_out/ncc_summary.txt 2> gpurun_out/ncc_summary.err; echo "summary rc=$?"
grep -n "=====\|duration\|issue slots\|LSU pipe\|occupancy\|registers\|warp instructions\|L2 hit\|FMA pipe\|XU" gpurun_out/ncc_summary.txt | head -60
