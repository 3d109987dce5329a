# usage (under gpurun): bash scripts/gpu_append.sh <tag> -- b2_paged_decode_append: parity tests, then the bench step fused vs two calls
tag=${1:-a}
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_append.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -12) > gpurun_out/append_pytest_$tag.log
cat gpurun_out/append_pytest_$tag.log
for wl in llama3-8b_b64_ctx4k_bf16kv llama3-70b_b256_ctx8k_bf16kv llama3-70b_tp8_shard; do
  for ap in 1 0 1 0; do
    B2_BENCH_APPEND=$ap timeout 200 python bench.py --steps 200 --warmup 20 --lean --no-cpu --workload $wl 2>/dev/null | tail -1 > gpurun_out/append_bench.json
    python - <<PY
import json
try:
    j=json.load(open('gpurun_out/append_bench.json')); r=j['roofline']
    print('$wl', 'append=$ap', 'value', j['value'], 'ms', j['ms_per_step'], 'kern', r['kernel_ms'], 'kern_min', r['kernel_ms_min'], 'e2e', j['e2e']['value'], 'launches', j['gpu_launches'], j['config'].get('num_splits'))
except Exception as e:
    print('$wl', 'append=$ap', 'ERR', e, open('gpurun_out/append_bench.json').read()[:300])
PY
  done
done 2>&1 | tee gpurun_out/append_ab_$tag.txt
