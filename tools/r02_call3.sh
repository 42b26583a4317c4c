#!/bin/bash
# Round 2, third GPU pass: full GPU suite (regression), headline bench with the pipe roofline, C5 bench both arms.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -x -m gpu --no-header -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "gpu suite rc=$?"
tail -n 6 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?"
cat gpurun_out/bench_ours.json; tail -n 5 gpurun_out/bench_ours.err
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 > gpurun_out/bench_c5_ours.json 2> gpurun_out/bench_c5_ours.err; echo "c5 ours rc=$?"
cat gpurun_out/bench_c5_ours.json; tail -n 5 gpurun_out/bench_c5_ours.err
timeout 600 python bench.py --workload c5 --impl reference --steps 5 --warmup 3 > gpurun_out/bench_c5_ref.json 2> gpurun_out/bench_c5_ref.err; echo "c5 ref rc=$?"
cat gpurun_out/bench_c5_ref.json; tail -n 5 gpurun_out/bench_c5_ref.err
