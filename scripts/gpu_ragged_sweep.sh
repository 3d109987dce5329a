# experiment: static share / chunk size of the dynamic schedule on ragged batches
for wl in llama3-8b_b64_ctx4k_bf16kv llama3-8b_b128_ctx8k_fp8kv; do
 for pct in 50 70 85 95; do for ct in 256 512; do
  B2_STATIC_PCT=$pct B2_CHUNK_TOKENS=$ct python bench.py --steps 200 --warmup 20 --no-cpu --lean --ragged --workload $wl --splits 0 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']
print('$wl pct $pct ct $ct kern_ms', r['kernel_ms'], 'ach', r['achieved'])"
 done; done
done
