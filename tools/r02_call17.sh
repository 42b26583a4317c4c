#!/bin/bash
set -u
mkdir -p gpurun_out
V=$PWD/geometry-grounded-gaussian-splatting_b200/diff_gaussian_rasterization/variants
timeout 900 python -m pytest tests/test_fullsize_parity_gpu.py tests/test_parity_gpu.py tests/test_edge_gpu.py -q -x --no-header -p no:cacheprovider > gpurun_out/pytest_cls.log 2>&1; echo "parity rc=$?"
tail -n 2 gpurun_out/pytest_cls.log
bash tools/variants.sh base cls0 cls1 cls5
bash tools/variants.sh --workload c4 base cls0
GGS_RASTERIZER_LIB=$V/libggs_rasterizer_stats.so timeout 600 python tools/walk_stats.py c3 > gpurun_out/walk_stats.json 2> gpurun_out/walk_stats.err; echo "walk stats rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/walk_stats.json'))['c3']
for k,v in d.items(): print(k, v['walked'], v['evaluated'], v['mask_words'])"
