TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
mkdir -p gpurun_out
for W in c3 c4; do for PULL in 0 1; do
GGS_PEER_PULL=$PULL timeout 300 $TR bench.py --gpus 8 --workload $W --steps 30 --warmup 5 > gpurun_out/bench_n8_${W}_pull${PULL}.json 2> gpurun_out/bench_n8_${W}_pull${PULL}.err; echo "$W pull=$PULL rc=$?"
python -c "import json; d=json.loads(open('gpurun_out/bench_n8_${W}_pull${PULL}.json').read().strip().splitlines()[-1]); print(d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],4), round(d['e2e']['value'],1), d['config']['parallelism'][-70:])"
done; done
