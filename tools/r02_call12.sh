#!/bin/bash
# Round 2, session 3: patch-NCC kernel without the scratch column (occupancy variants), C5 bench with later rounds.
set -u
mkdir -p gpurun_out
V=$PWD/geometry-grounded-gaussian-splatting_b200/warp_patch_ncc/variants
timeout 600 python -m pytest tests/test_patch_ncc_gpu.py -q -x --no-header -p no:cacheprovider > gpurun_out/pytest_ncc.log 2>&1; echo "ncc tests rc=$?"; tail -n 2 gpurun_out/pytest_ncc.log
for b in base b5 b6 b8; do
  if [ "$b" = "base" ]; then unset GGS_PATCH_NCC_LIB; else export GGS_PATCH_NCC_LIB=$V/libggs_patch_ncc_$b.so; fi
  timeout 300 python tools/ncc_bench.py > gpurun_out/ncc_bench_$b.json 2> gpurun_out/ncc_bench_$b.err; echo "ncc $b rc=$?"
  python -c "import json; d=json.load(open('gpurun_out/ncc_bench_$b.json')); print('$b', {o: (round(d[o]['ours_ms'],4), round(d[o].get('reference_ms',0),4), d[o].get('bit_equal')) for o in ('raster','random')})"
done
unset GGS_PATCH_NCC_LIB
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_ours.json 2> gpurun_out/bench_c5_ours.err; echo "c5 rc=$?"
python -c "import json; d=json.loads(open('gpurun_out/bench_c5_ours.json').read().strip().splitlines()[-1]); print('c5', round(d['value'],1), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'later', d.get('later_rounds'))"
