#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fullsize_parity_gpu.py tests/test_parity_gpu.py tests/test_edge_gpu.py tests/test_sort_gpu.py tests/test_integrate_gpu.py tests/test_reference_wrapper_gpu.py -q -x --no-header -p no:cacheprovider > gpurun_out/pytest_k64.log 2>&1; echo "parity rc=$?"
tail -n 2 gpurun_out/pytest_k64.log
bash tools/variants.sh base bwd4
bash tools/variants.sh --workload c4 base
