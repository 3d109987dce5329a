# dynamic-schedule parameters on one rank's share of BASELINE configs[3] at tp = 8 (256 items on 296 CTA slots)
for s in 1 0; do
  python bench.py --steps 200 --warmup 20 --no-cpu --workload llama3-70b_tp8_shard --splits $s | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('splits $s default', 'kern_ms', r['kernel_ms'], 'ach', r['achieved'])"
done
for pct in 50 70 85 95; do for ct in 128 256 512; do
  B2_STATIC_PCT=$pct B2_CHUNK_TOKENS=$ct python bench.py --steps 200 --warmup 20 --no-cpu --workload llama3-70b_tp8_shard --splits 0 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('pct $pct ct $ct', 'kern_ms', r['kernel_ms'], 'ach', r['achieved'])"
done; done
for sp in 2 4; do
  python bench.py --steps 200 --warmup 20 --no-cpu --workload llama3-70b_tp8_shard --splits $sp | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('splits $sp', 'kern_ms', r['kernel_ms'], 'ach', r['achieved'])"
done
