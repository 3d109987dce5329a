# usage: bash scripts/gpu_scale.sh <N> <tag>   (under gpurun --gpus N): headline weak scaling + BASELINE configs[3] KV-head sharded (strong)
N=$1; tag=${2:-x}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" 2>gpurun_out/scale_${tag}_n$N.err | tail -1; }
run --steps 300 --warmup 20 --no-cpu > gpurun_out/scale_${tag}_n${N}_weak.json
run --steps 200 --warmup 20 --no-cpu --workload llama3-70b_b256_ctx8k_bf16kv --strong > gpurun_out/scale_${tag}_n${N}_config4_strong.json
run --steps 200 --warmup 20 --no-cpu --workload llama3-8b_b128_ctx8k_fp8kv > gpurun_out/scale_${tag}_n${N}_fp8_weak.json
cat gpurun_out/scale_${tag}_n${N}_*.json | python -c "
import json,sys
for l in sys.stdin:
    try: j=json.loads(l)
    except Exception: print('ERR', l[:200]); continue
    print(j['n_gpus'], j['config']['workload'], j['scaling'], 'value', j['value'], 'ms', j['ms_per_step'], j['config'].get('parallelism'), j.get('clocks',{}).get('reasons'))"
