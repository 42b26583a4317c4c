#!/bin/bash
set -u
mkdir -p gpurun_out
V=$PWD/geometry-grounded-gaussian-splatting_b200/diff_gaussian_rasterization/variants
for P in 1000000 2000000; do
for n in base mv128x6 mv128x8 mv256x4 mv256x3 mv64x8 mv64x12; do
  if [ "$n" = "base" ]; then unset GGS_RASTERIZER_LIB; else export GGS_RASTERIZER_LIB=$V/libggs_rasterizer_$n.so; fi
  timeout 120 python tools/sh_sum_bench.py $P 8 2>&1 | grep "^{" | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$n', d['P'], round(d['us'],1), 'us', round(d['frac_hbm'],3))"
done; done
