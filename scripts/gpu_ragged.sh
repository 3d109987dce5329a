# experiment: ragged batches (log-uniform 128..ctx), single-pass grid vs dynamic schedule vs host planner
for wl in llama3-8b_b64_ctx4k_bf16kv llama3-8b_b128_ctx8k_fp8kv llama3-70b_b256_ctx8k_bf16kv; do
  for s in 1 0 -1; do
    python bench.py --steps 300 --warmup 20 --no-cpu --lean --ragged --workload $wl --splits $s | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; c=j['config']
print(c['workload'], 'asked', $s, 'ran', c['num_splits'], 'bytes', c['algorithmic_bytes_per_step_per_gpu'], 'kern_ms', r['kernel_ms'], 'ach', r['achieved'], 'e2e', j['e2e']['value'])"
  done
done
