# experiment: fp8 decode, 2 vs 3 CTAs per SM
(timeout 900 python -m pytest tests/test_gpu_parity.py -q -x 2>&1 | tail -3)
for c in 3 2; do
  for s in 0 1; do
    B2_FP8_CTAS=$c timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu --workload llama3-8b_b128_ctx8k_fp8kv --splits $s 2>&1 | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; c=j['config']
print('ctas $c', c['workload'],'bs',c['block_size'],'splits',c['num_splits'],'kern_ms',r['kernel_ms'],'ach',r['achieved'])"
  done
done
