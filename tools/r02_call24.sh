#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_integrate_gpu.py tests/test_points_golden.py tests/test_sample_depth_gpu.py tests/test_reference_callers_gpu.py tests/test_parity_gpu.py -q -x --no-header -p no:cacheprovider > gpurun_out/pytest_q.log 2>&1; echo "tests rc=$?"
tail -n 2 gpurun_out/pytest_q.log
timeout 900 python tools/points_bench.py --what integrate,sample_depth,evaluate_sdf --scene c5 --sample-scene c3 --iters 5 --out gpurun_out/points_bench.json > gpurun_out/points_bench.log 2>&1; echo "points_bench rc=$?"
grep -o '"ours_fwd_ms": [0-9.]*\|"ours_ms": [0-9.]*\|"ref_ms": [0-9.]*\|bit_equal": [a-z]*\|"speedup[a-z_]*": [0-9.]*' gpurun_out/points_bench.log | tr '\n' ' '; echo
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_ours.json 2> gpurun_out/bench_c5_ours.err; echo "c5 ours rc=$?"
python -c "
import json; o=json.loads(open('gpurun_out/bench_c5_ours.json').read().strip().splitlines()[-1]); print('c5', round(o['value'],1), 'e2e', round(o['e2e']['value'],1), 'later', round(o['later_rounds']['value'],1), 'kernel ms', round(o['roofline']['ms'],3), 'frac', round(o['roofline']['frac'],3))"
