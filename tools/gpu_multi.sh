#!/bin/bash
# Multi-GPU pass (gpurun --gpus N): peer-exchange check against NCCL, then bench at N ranks with and without it.
set -u
N=${N:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
if [ "${PYTEST:-0}" = "1" ]; then
  timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
fi
[ "${C1:-1}" = "1" ] && PEER_WORKLOAD=c1 timeout 240 $TR tools/peer_check.py > gpurun_out/peer_check_c1.log 2>&1; echo "peer_check c1 rc=$?"
grep "rank\|nccl\|Error\|error" gpurun_out/peer_check_c1.log | grep -v Warning | tail -n 14
PEER_WORKLOAD=c3 timeout 300 $TR tools/peer_check.py > gpurun_out/peer_check_c3.log 2>&1; echo "peer_check c3 rc=$?"
grep "rank\|nccl\|Error\|error" gpurun_out/peer_check_c3.log | grep -v Warning | tail -n 14
if [ "${BENCH:-1}" = "1" ]; then
timeout 400 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n${N}_peer.json 2> gpurun_out/bench_n${N}_peer.err; echo "bench peer rc=$?"
python -c "import json,sys; d=json.loads(open('gpurun_out/bench_n${N}_peer.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['parallelism'][-60:])"
GGS_PEER=0 timeout 400 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n${N}_nccl.json 2> gpurun_out/bench_n${N}_nccl.err; echo "bench nccl rc=$?"
python -c "import json,sys; d=json.loads(open('gpurun_out/bench_n${N}_nccl.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['parallelism'][-60:])"
[ "${N1:-1}" = "1" ] && timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench n1 rc=$?"
python -c "import json,sys; d=json.load(open('gpurun_out/bench_n1.json')); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'])"
fi
if [ "${C4:-0}" = "1" ]; then
timeout 400 $TR bench.py --gpus $N --workload c4 --steps 30 --warmup 5 > gpurun_out/bench_n${N}_c4_peer.json 2> gpurun_out/bench_n${N}_c4_peer.err; echo "bench c4 peer rc=$?"
python -c "import json,sys; d=json.loads(open('gpurun_out/bench_n${N}_c4_peer.json').read().strip().splitlines()[-1]); print('c4', d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['parallelism'][-60:])"
fi
