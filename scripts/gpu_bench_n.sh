# usage: bash scripts/gpu_bench_n.sh <tag> <ngpus> [steps]   (under gpurun --gpus N): the driver's N-GPU bench command
tag=${1:-bn}; n=${2:-2}; k=${3:-20}
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $n --steps $k --warmup 5 > gpurun_out/bench_${tag}_n$n.json 2> gpurun_out/bench_${tag}_n$n.err
echo "bench rc=$?"; tail -c 800 gpurun_out/bench_${tag}_n$n.err
python - <<PY
import json
j=json.loads([l for l in open('gpurun_out/bench_${tag}_n$n.json') if l.startswith('{')][-1])
print('value',j['value'],'ms',j['ms_per_step'],'e2e',j['e2e']['value'])
for k in ('strong_config4','sharded_parity','tp_serving'):
    print(k, json.dumps(j.get(k))[:1800])
PY
