# usage (under gpurun): bash scripts/gpu_append_final.sh -- closing run with the fused append step: whole GPU suite, the driver-style
# bench line, the reference arm, then (each only after its plain command exited 0) the ncu launch list and one full capture
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/af_pytest.log; cat gpurun_out/af_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 500 python bench.py --steps 20 --warmup 5 2>gpurun_out/af_bench.err | tail -1 > gpurun_out/af_bench_n1.json; echo "bench rc=$?"
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 > gpurun_out/af_bench_reference.json
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --lean"
$CMD > gpurun_out/af_plain.log 2>&1 || { tail -5 gpurun_out/af_plain.log; exit 1; }
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_decode|reshape_and_cache|merge_|zero_counter' -c 400 --csv \
    --log-file gpurun_out/launches_af.csv $CMD > gpurun_out/af_ncu_launch.log 2>&1
C2="python bench.py --steps 8 --warmup 3 --no-cpu --lean"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'paged_decode_split_kernel' -s 6 -c 1 -f \
    -o gpurun_out/prof_af_append $C2 > gpurun_out/af_ncu_full.log 2>&1
python - <<'PY'
import json
j=json.load(open('gpurun_out/af_bench_n1.json'))
print('value', j['value'], 'ms', j['ms_per_step'], 'kern', j['roofline']['kernel_ms'], j['roofline']['frac'], 'e2e', j['e2e']['value'], 'launches', j['gpu_launches'])
print(j['config'].get('step_impl'))
for k,v in j['configs'].items(): print(k, v.get('value'), v.get('kernel_ms'), v.get('frac'), v.get('step_impl','')[:40])
print('decode_tok_s', j['decode_tok_s']['value'])
PY
ls -la gpurun_out | tail -8
