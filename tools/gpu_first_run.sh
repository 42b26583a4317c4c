#!/bin/bash
# First GPU pass over everything written without hardware at the end of round 1.
# Run under gpurun on ONE GPU:   gpurun --timeout 1500 -- 'bash tools/gpu_first_run.sh'
# Each step is bounded by its own timeout and writes to gpurun_out/; nothing here runs under a profiler except the
# two ncu captures at the end (NCU=1), each after its plain command has exited 0.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt
# 1. the measured parity suite first (everything that ran on hardware before), then the first-run tests one file at a time
timeout 1200 python -m pytest tests -q -m "gpu and not first_hardware_run" > gpurun_out/pytest_measured.log 2>&1; echo "measured suite rc=$?"
tail -n 3 gpurun_out/pytest_measured.log
for f in test_ssim_gpu test_sample_depth_gpu test_patch_ncc_gpu; do
  timeout 900 python -m pytest tests/$f.py -q -m "gpu and first_hardware_run" -s > gpurun_out/$f.log 2>&1; echo "$f rc=$?"
  tail -n 5 gpurun_out/$f.log
done
# 2. numbers: loss kernels against their references, then a training iteration through every kernel
timeout 600 python tools/ncc_bench.py > gpurun_out/ncc_bench.json 2> gpurun_out/ncc_bench.err; echo "ncc_bench rc=$?"; cat gpurun_out/ncc_bench.json
timeout 600 python tools/ssim_bench.py > gpurun_out/ssim_bench.json 2> gpurun_out/ssim_bench.err; echo "ssim_bench rc=$?"; cat gpurun_out/ssim_bench.json
timeout 600 python tools/train_step_smoke.py > gpurun_out/train_smoke.json 2> gpurun_out/train_smoke.err; echo "train_smoke rc=$?"; cat gpurun_out/train_smoke.json
# 3. the headline bench, unchanged path (sanity that nothing regressed)
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?"
cat gpurun_out/bench_ours.json
if [ "${NCU:-0}" = "1" ]; then
  timeout 900 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:ggs::ncc -c 2 \
      -f -o gpurun_out/ncc_full python tools/ncc_bench.py --iters 2 > gpurun_out/ncu_ncc.log 2>&1; echo "ncu ncc rc=$?"
  timeout 900 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:ggs::ssim -c 6 \
      -f -o gpurun_out/ssim_full python tools/ssim_bench.py --iters 2 > gpurun_out/ncu_ssim.log 2>&1; echo "ncu ssim rc=$?"
fi
