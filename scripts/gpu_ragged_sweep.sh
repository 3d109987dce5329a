# experiment: static share / chunk size of the dynamic schedule on ragged batches (and the uniform fp8 config, which
# also takes the dynamic schedule)
for wl in "llama3-8b_b64_ctx4k_bf16kv --ragged" "llama3-8b_b128_ctx8k_fp8kv --ragged" "llama3-70b_b256_ctx8k_bf16kv --ragged" "llama3-8b_b128_ctx8k_fp8kv"; do
 for pct in 85 92 97; do for ct in 64 128 256 512; do
  B2_STATIC_PCT=$pct B2_CHUNK_TOKENS=$ct python bench.py --steps 200 --warmup 20 --no-cpu --lean --workload $wl --splits 0 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']
print('$wl pct $pct ct $ct kern_ms', r['kernel_ms'], 'ach', r['achieved'])"
 done; done
done
