#!/bin/bash
# Round 2, first GPU pass (one GPU): parity at the quoted sizes, point-query baselines, loss-kernel benches, sanitizer.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt
timeout 1500 python -m pytest tests/test_fullsize_parity_gpu.py tests/test_edge_gpu.py -q -x --no-header -p no:cacheprovider > gpurun_out/pytest_fullsize.log 2>&1; echo "fullsize rc=$?"
tail -n 15 gpurun_out/pytest_fullsize.log
timeout 900 python tools/points_bench.py --what integrate,sample_depth,evaluate_sdf --scene c5 --sample-scene c2+c3 --iters 5 --out gpurun_out/points_bench_r02_v0.json > gpurun_out/points_bench.log 2>&1; echo "points_bench rc=$?"
tail -n 8 gpurun_out/points_bench.log
timeout 600 python tools/ncc_bench.py > gpurun_out/ncc_bench.json 2> gpurun_out/ncc_bench.err; echo "ncc_bench rc=$?"; cat gpurun_out/ncc_bench.json
timeout 600 python tools/ssim_bench.py > gpurun_out/ssim_bench.json 2> gpurun_out/ssim_bench.err; echo "ssim_bench rc=$?"; cat gpurun_out/ssim_bench.json
timeout 600 python -m pytest tests/test_patch_ncc_gpu.py tests/test_ssim_gpu.py -q --no-header -p no:cacheprovider > gpurun_out/pytest_n4.log 2>&1; echo "n4 tests rc=$?"; tail -n 3 gpurun_out/pytest_n4.log
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?"
cat gpurun_out/bench_ours.json
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_run.py > gpurun_out/sanitizer_$tool.log 2>&1; echo "$tool rc=$?"
  tail -n 6 gpurun_out/sanitizer_$tool.log
done
