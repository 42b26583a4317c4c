#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fullsize_parity_gpu.py tests/test_parity_gpu.py tests/test_edge_gpu.py tests/test_sort_gpu.py tests/test_integrate_gpu.py -q -x --no-header -p no:cacheprovider > gpurun_out/pytest_s32.log 2>&1; echo "parity rc=$?"
tail -n 2 gpurun_out/pytest_s32.log
bash tools/variants.sh base prev
