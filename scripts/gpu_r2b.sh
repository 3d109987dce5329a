# usage: bash scripts/gpu_r2b.sh <tag>  (under gpurun): GPU tests, driver-style bench line with all sub-records, gather ceiling, vLLM fixture
tag=${1:-r2b}
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -40) > gpurun_out/pytest_$tag.log 2>&1
tail -5 gpurun_out/pytest_$tag.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_${tag}_k20.json 2> gpurun_out/bench_${tag}_k20.err
echo "bench k20 rc=$?"; tail -c 600 gpurun_out/bench_${tag}_k20.err
timeout 600 python bench.py --lean --steps 500 --warmup 20 > gpurun_out/bench_${tag}_k500.json 2> gpurun_out/bench_${tag}_k500.err
echo "bench k500 rc=$?"
timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_${tag}_ref.json 2>&1
timeout 120 scripts/bin/gather_bw > gpurun_out/gather_bw_$tag.txt 2>&1; cat gpurun_out/gather_bw_$tag.txt
timeout 120 scripts/bin/gather_bw 256 8 8192 > gpurun_out/gather_bw_${tag}_cfg4.txt 2>&1; cat gpurun_out/gather_bw_${tag}_cfg4.txt
timeout 600 python tests/golden/make_vllm_crosscheck.py > gpurun_out/make_vllm_$tag.log 2>&1; tail -3 gpurun_out/make_vllm_$tag.log
python - <<PY
import json
for f in ("gpurun_out/bench_${tag}_k20.json","gpurun_out/bench_${tag}_k500.json"):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f,'ERR',e); continue
    r=j['roofline']
    print(f, 'value',j['value'],'kern_ms',r['kernel_ms'],'frac',r['frac'],'e2e',j['e2e']['value'],j['e2e']['ms_per_step'],'h2d',j['e2e']['h2d_bytes_per_step'])
    for k,v in j.get('configs',{}).items(): print('  ',k,{x:v.get(x) for x in ('value','kernel_ms','achieved','frac','num_splits','error')})
    if 'decode_tok_s' in j: print('   decode_tok_s', {x:j['decode_tok_s'].get(x) for x in ('value','device_tok_s','frac_of_roofline','error')})
    for c in j.get('comparator',{}).get('results',[]): print('   cmp', c.get('name'), c.get('workload'), c.get('kernel_ms'), c.get('GBps'), c.get('max_abs_diff_vs_ours'), c.get('unavailable'))
    if 'comparator' in j and 'results' not in j['comparator']: print('   cmp', j['comparator'])
PY
