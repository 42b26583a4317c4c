#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fullsize_parity_gpu.py tests/test_parity_gpu.py tests/test_edge_gpu.py tests/test_multiview_gpu.py -q -x --no-header -p no:cacheprovider > gpurun_out/pytest_bw.log 2>&1; echo "parity rc=$?"
tail -n 2 gpurun_out/pytest_bw.log
bash tools/variants.sh base prev
bash tools/variants.sh --workload c4 base prev
