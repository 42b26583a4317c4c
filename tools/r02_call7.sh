#!/bin/bash
# Round 2, session 3: quiet-rule thinning / folding of the forward's walk (separate FAR-rule export for the backward).
set -u
mkdir -p gpurun_out
V=$PWD/geometry-grounded-gaussian-splatting_b200/diff_gaussian_rasterization/variants
timeout 900 python -m pytest tests/test_fullsize_parity_gpu.py tests/test_parity_gpu.py tests/test_sample_depth_gpu.py tests/test_integrate_gpu.py tests/test_points_golden.py tests/test_edge_gpu.py -q -x --no-header -p no:cacheprovider -rs > gpurun_out/pytest_walk.log 2>&1; echo "parity rc=$?"
tail -n 5 gpurun_out/pytest_walk.log
bash tools/variants.sh base
GGS_RASTERIZER_LIB=$V/libggs_rasterizer_stats.so timeout 600 python tools/walk_stats.py c3 > gpurun_out/walk_stats.json 2> gpurun_out/walk_stats.err; echo "walk stats rc=$?"
cat gpurun_out/walk_stats.json | tr -d '\n ' ; echo
timeout 900 python tools/points_bench.py --what sample_depth,evaluate_sdf --scene c5 --sample-scene c2+c3 --iters 5 --out gpurun_out/points_bench_r02_v3.json > gpurun_out/points_bench.log 2>&1; echo "points_bench rc=$?"
grep -o '"ours_fwd_ms": [0-9.]*\|"ours_ms": [0-9.]*\|bit_equal": [a-z]*\|"speedup[a-z_]*": [0-9.]*' gpurun_out/points_bench.log | tr '\n' ' '; echo
