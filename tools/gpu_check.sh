#!/bin/bash
# One GPU-box pass: golden vectors, GPU tests, bench (both arms), ncu launch list.  Run under gpurun.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt
if [ "${GOLDEN:-0}" = "1" ]; then
  timeout 300 python tests/golden/make_golden.py > gpurun_out/golden.log 2>&1; echo "golden rc=$?"
fi
timeout 1200 python -m pytest tests -q -m gpu ${PYTEST_ARGS:-} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n ${TAIL:-40} gpurun_out/pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 30 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
cat gpurun_out/bench_ours.json
if [ "${NCU:-0}" = "1" ]; then
  timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
  echo "ncu rc=$?"
fi
if [ "${FULL:-0}" = "1" ]; then
  # one step's worth of our kernels with the full metric set (DRAM traffic, pipe utilisation, source pages)
  timeout 1200 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:ggs:: --launch-skip ${FULL_SKIP:-120} -c ${FULL_COUNT:-24} \
      -f -o gpurun_out/full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
