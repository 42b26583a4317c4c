#!/bin/bash
# Full single-GPU validation: the whole GPU suite, both bench arms, C5 both arms, point queries, launch list + full ncu.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt
timeout 1800 python -m pytest tests -q -m gpu --no-header -p no:cacheprovider -rs > gpurun_out/pytest_gpu.log 2>&1; echo "gpu suite rc=$?"
tail -n 12 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 30 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
timeout 600 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
python - <<'PY'
import json
o = json.loads(open("gpurun_out/bench_ours.json").read().strip().splitlines()[-1])
r = json.loads(open("gpurun_out/bench_ref.json").read().strip().splitlines()[-1])
print("ours", round(o["value"], 1), "ms", round(o["ms_per_step"], 4), "e2e", round(o["e2e"]["value"], 1), "| ref", round(r["value"], 1), "e2e", round(r["e2e"]["value"], 1),
      "| ratio", round(o["value"] / r["value"], 3), "e2e ratio", round(o["e2e"]["value"] / r["e2e"]["value"], 3))
print({k: round(v["ms"], 4) for k, v in o["stages"].items()})
print("roofline", {k: (round(v, 3) if isinstance(v, float) else v) for k, v in o["roofline"].items() if k in ("kernel", "bound", "frac", "frac_xu", "frac_fp32")}, "clocks", o["clocks"])
PY
if [ "${C5:-1}" = "1" ]; then
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 > gpurun_out/bench_c5_ours.json 2> gpurun_out/bench_c5_ours.err; echo "c5 ours rc=$?"
timeout 600 python bench.py --workload c5 --impl reference --steps 5 --warmup 3 > gpurun_out/bench_c5_ref.json 2> gpurun_out/bench_c5_ref.err; echo "c5 ref rc=$?"
python - <<'PY'
import json
o = json.loads(open("gpurun_out/bench_c5_ours.json").read().strip().splitlines()[-1])
r = json.loads(open("gpurun_out/bench_c5_ref.json").read().strip().splitlines()[-1])
print("c5 ours", round(o["value"], 1), "e2e", round(o["e2e"]["value"], 1), "later", round(o["later_rounds"]["value"], 1), "| ref", round(r["value"], 1), "e2e", round(r["e2e"]["value"], 1), "| ratio", round(o["value"] / r["value"], 3))
PY
timeout 900 python tools/points_bench.py --what integrate,sample_depth,evaluate_sdf --scene c5 --sample-scene c2+c3 --iters 5 --out gpurun_out/points_bench.json > gpurun_out/points_bench.log 2>&1; echo "points_bench rc=$?"
grep -o '"ours_fwd_ms": [0-9.]*\|"ours_ms": [0-9.]*\|bit_equal": [a-z]*\|"speedup[a-z_]*": [0-9.]*' gpurun_out/points_bench.log | tr '\n' ' '; echo
fi
if [ "${NCU:-1}" = "1" ]; then
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
echo "ncu launches rc=$?"
timeout 1200 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:ggs:: --launch-skip ${FULL_SKIP:-110} -c ${FULL_COUNT:-22} \
    -f -o gpurun_out/full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
python tools/ncu_summary.py gpurun_out/full.ncu-rep > gpurun_out/full_summary.txt 2> gpurun_out/full_summary.err; echo "summary rc=$?"
grep "^=====" gpurun_out/full_summary.txt | sort | uniq -c
fi
