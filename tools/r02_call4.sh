#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_reference_wrapper_gpu.py tests/test_reference_callers_gpu.py -q --no-header -p no:cacheprovider > gpurun_out/pytest_refapp.log 2>&1; echo "refapp rc=$?"
tail -n 40 gpurun_out/pytest_refapp.log
