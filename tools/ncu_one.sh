#!/bin/bash
# ncu --set full of a few launches of ONE kernel (regex in $1, report name in $2), summarised.  After a plain run.
set -u
mkdir -p gpurun_out
K=${1:-render_forward}; OUT=${2:-one}
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -n 3 gpurun_out/plain.log; exit 1; }
timeout 900 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:$K --launch-skip ${SKIP:-12} -c ${COUNT:-1} \
    -f -o gpurun_out/$OUT python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"
python tools/ncu_summary.py gpurun_out/$OUT.ncu-rep --lines 60 > gpurun_out/${OUT}_summary.txt 2> gpurun_out/${OUT}_summary.err; echo "summary rc=$?"
head -n 24 gpurun_out/${OUT}_summary.txt
