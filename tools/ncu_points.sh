#!/bin/bash
# ncu --set full of the point-query kernels inside tools/points_bench.py (integrate on c5).
set -u
mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:"integrate_kernel|sample_forward_kernel" --launch-skip ${SKIP:-2} -c ${COUNT:-2} \
    -f -o gpurun_out/points python tools/points_bench.py --what ${WHAT:-integrate} --scene c5 --sample-scene c3 --iters 2 --out gpurun_out/points_ncu.json > gpurun_out/ncu_points.log 2>&1
echo "ncu rc=$?"
python tools/ncu_summary.py gpurun_out/points.ncu-rep --lines 45 > gpurun_out/points_summary.txt 2> gpurun_out/points_summary.err; echo "summary rc=$?"
head -n 70 gpurun_out/points_summary.txt
