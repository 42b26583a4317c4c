#!/bin/bash
# Round 2, session 3, first GPU pass (one GPU): A/B of the quiet-splat shortcut, regression of the whole GPU suite,
# both bench arms, ncu launch list + full capture, compute-sanitizer.
set -u
mkdir -p gpurun_out
V=geometry-grounded-gaussian-splatting_b200/diff_gaussian_rasterization/variants
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
cat gpurun_out/bench_ours.json
GGS_RASTERIZER_LIB=$PWD/$V/libggs_rasterizer_quiet16.so timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_quiet16.json 2> gpurun_out/bench_quiet16.err; echo "bench quiet16 rc=$?"
python - <<'PY'
import json
for n in ("ours", "quiet16"):
    try:
        d = json.loads(open(f"gpurun_out/bench_{n}.json").read().strip().splitlines()[-1])
        print(n, "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"], 1), {k: round(v["ms"], 4) for k, v in d["stages"].items()})
    except Exception as ex:
        print(n, "unreadable", ex)
PY
timeout 600 python bench.py --impl reference --steps 30 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 900 python tools/points_bench.py --what integrate,sample_depth,evaluate_sdf --scene c5 --sample-scene c2+c3 --iters 5 --out gpurun_out/points_bench_r02_v2.json > gpurun_out/points_bench.log 2>&1; echo "points_bench rc=$?"
tail -n 4 gpurun_out/points_bench.log
timeout 1800 python -m pytest tests -q -x -m gpu --no-header -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "gpu suite rc=$?"
tail -n 8 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
echo "ncu launches rc=$?"
timeout 1200 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:ggs:: --launch-skip ${FULL_SKIP:-120} -c ${FULL_COUNT:-24} \
    -f -o gpurun_out/full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
python tools/ncu_summary.py gpurun_out/full.ncu-rep > gpurun_out/full_summary.txt 2> gpurun_out/full_summary.err; echo "summary rc=$?"
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_run.py --small > gpurun_out/sanitizer_$tool.log 2>&1; echo "$tool rc=$?"
  tail -n 6 gpurun_out/sanitizer_$tool.log
done
