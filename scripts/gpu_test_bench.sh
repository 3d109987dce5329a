# usage: bash scripts/gpu_test_bench.sh <tag>   (run under gpurun): GPU parity tests, then one bench line per workload/schedule
tag=${1:-x}
(timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -40) > gpurun_out/pytest_$tag.log 2>&1
tail -8 gpurun_out/pytest_$tag.log
rm -f gpurun_out/sweep_$tag.jsonl
for wl in llama3-8b_b64_ctx4k_bf16kv llama3-8b_b128_ctx8k_fp8kv llama3-70b_b256_ctx8k_bf16kv oracle_b4_ctx512; do
  for s in -1 0 1; do
    timeout 300 python bench.py --steps 300 --warmup 20 --no-cpu --workload $wl --splits $s 2>&1 | tail -1 >> gpurun_out/sweep_$tag.jsonl
  done
done
python - <<PY
import json
for l in open('gpurun_out/sweep_$tag.jsonl'):
    try: j=json.loads(l)
    except Exception: print('ERR', l[:300]); continue
    c=j['config']; r=j['roofline']
    print(c['workload'],'splits',c['num_splits'],'value',j['value'],'pct8T',j['pct_of_8TBps'],'kern_ms',r['kernel_ms'],'ach',r['achieved'],'frac',r['frac'],'e2e',j['e2e']['value'],'e2e_ms',j['e2e']['ms_per_step'],j['clocks']['reasons'])
PY
