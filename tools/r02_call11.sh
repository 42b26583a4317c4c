#!/bin/bash
# Round 2, session 3: forward without the blocking read-back of the pair count (capacity hints), both ways.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_edge_gpu.py tests/test_parity_gpu.py tests/test_fullsize_parity_gpu.py tests/test_reference_wrapper_gpu.py tests/test_integrate_gpu.py tests/test_sample_depth_gpu.py tests/test_multiview_gpu.py -q -x --no-header -p no:cacheprovider -rs > gpurun_out/pytest_hint.log 2>&1; echo "tests rc=$?"
tail -n 4 gpurun_out/pytest_hint.log
for H in 0 1; do
  GGS_CAPACITY_HINTS=$H timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_hint$H.json 2> gpurun_out/bench_hint$H.err; echo "hints=$H rc=$?"
  python -c "import json; d=json.loads(open('gpurun_out/bench_hint$H.json').read().strip().splitlines()[-1]); print('hints=$H', 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1))"
done
