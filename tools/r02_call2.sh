#!/bin/bash
# Round 2, second GPU pass: the regrouped point-query kernels (parity first, then times against the reference).
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_integrate_gpu.py tests/test_sample_depth_gpu.py tests/test_points_golden.py tests/test_edge_gpu.py -q -x --no-header -p no:cacheprovider > gpurun_out/pytest_points.log 2>&1; echo "points tests rc=$?"
tail -n 25 gpurun_out/pytest_points.log
timeout 900 python tools/points_bench.py --what integrate,sample_depth,evaluate_sdf --scene c5 --sample-scene c2+c3 --iters 5 --out gpurun_out/points_bench_r02_v1.json > gpurun_out/points_bench.log 2>&1; echo "points_bench rc=$?"
tail -n 8 gpurun_out/points_bench.log
