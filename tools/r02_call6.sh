#!/bin/bash
# Round 2, session 3, second GPU pass: walk statistics of the forward's bisection, A/B of kernel variants.
set -u
mkdir -p gpurun_out
V=$PWD/geometry-grounded-gaussian-splatting_b200/diff_gaussian_rasterization/variants
GGS_RASTERIZER_LIB=$V/libggs_rasterizer_stats.so timeout 600 python tools/walk_stats.py c3 c2 c4 > gpurun_out/walk_stats.json 2> gpurun_out/walk_stats.err; echo "walk stats rc=$?"
cat gpurun_out/walk_stats.json | tr -d '\n ' ; echo
bash tools/variants.sh base window occ9 occ10 look16 look32 hist296 base
GGS_RASTERIZER_LIB=$V/libggs_rasterizer_window.so timeout 900 python -m pytest tests/test_fullsize_parity_gpu.py tests/test_parity_gpu.py -q -x --no-header -p no:cacheprovider -rs > gpurun_out/pytest_window.log 2>&1; echo "window parity rc=$?"
tail -n 5 gpurun_out/pytest_window.log
