# usage: bash scripts/gpu_dyn_sweep.sh <tag>  -- parity first, then the schedule A/B on the headline config
tag=${1:-x}
(timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -30) > gpurun_out/pytest_$tag.log 2>&1
tail -5 gpurun_out/pytest_$tag.log
run() { # label, env..., args
  label=$1; shift
  line=$(env "$@" 2>/dev/null | tail -1)
  python - "$label" "$line" <<'PY'
import json,sys
try:
    j=json.loads(sys.argv[2]); r=j['roofline']
    print(sys.argv[1], 'value',j['value'],'kern_ms',r['kernel_ms'],'min',r['kernel_ms_min'],'ach',r['achieved'],'e2e',j['e2e']['value'], j['clocks']['reasons'])
except Exception as e:
    print(sys.argv[1], 'ERR', sys.argv[2][:200])
PY
}
B="timeout 300 python bench.py --steps 300 --warmup 20 --no-cpu"

run "c2 splits1     " X=1 $B --splits 1
for pct in 0 50 70 85; do for ct in 128 256 512; do
run "c2 dyn pct$pct ct$ct" B2_STATIC_PCT=$pct B2_CHUNK_TOKENS=$ct $B
done; done
run "c3 dyn default " X=1 $B --workload llama3-8b_b128_ctx8k_fp8kv
run "c3 dyn pct50   " B2_STATIC_PCT=50 $B --workload llama3-8b_b128_ctx8k_fp8kv
run "c3 splits1     " X=1 $B --workload llama3-8b_b128_ctx8k_fp8kv --splits 1
run "c4 dyn default " X=1 $B --workload llama3-70b_b256_ctx8k_bf16kv --steps 100
run "c4 splits1     " X=1 $B --workload llama3-70b_b256_ctx8k_bf16kv --steps 100 --splits 1
run "c1 dyn default " X=1 $B --workload oracle_b4_ctx512
run "c1 splits1     " X=1 $B --workload oracle_b4_ctx512 --splits 1
