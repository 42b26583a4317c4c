#!/bin/bash
# ncu --set full of the fused-SSIM kernels inside tools/ssim_bench.py.
set -u
mkdir -p gpurun_out
timeout 300 python tools/ssim_bench.py > gpurun_out/ssim_bench.json 2> gpurun_out/ssim_bench.err; echo "ssim bench rc=$?"; cat gpurun_out/ssim_bench.json
timeout 600 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:"ssim" --launch-skip 6 -c 3 \
    -f -o gpurun_out/ssim python tools/ssim_bench.py > gpurun_out/ncu_ssim.log 2>&1
echo "ncu rc=$?"
python tools/ncu_summary.py gpurun_out/ssim.ncu-rep --lines 30 > gpurun_out/ssim_summary.txt 2> gpurun_out/ssim_summary.err; echo "summary rc=$?"
grep -v "^       \|^      [0-9]" gpurun_out/ssim_summary.txt | head -80
