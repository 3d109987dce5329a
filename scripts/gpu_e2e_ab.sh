# usage (under gpurun): bash scripts/gpu_e2e_ab.sh -- DecodeStep with the fused plan call vs the two-call step, alternating on one box
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_append.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -6) > gpurun_out/e2e_ab_pytest.log; cat gpurun_out/e2e_ab_pytest.log
for i in 1 2 3; do for f in 1 0; do
B2_STEP_FUSE_APPEND=$f timeout 200 python bench.py --steps 300 --warmup 20 --lean --no-cpu 2>/dev/null | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print('fuse_append=$f', 'value', j['value'], 'e2e', j['e2e']['value'], 'e2e_ms', j['e2e']['ms_per_step'])"
done; done 2>&1 | tee gpurun_out/e2e_ab.txt
for wl in llama3-8b_b128_ctx8k_fp8kv llama3-70b_b256_ctx8k_bf16kv; do for f in 1 0; do
B2_STEP_FUSE_APPEND=$f timeout 200 python bench.py --steps 100 --warmup 10 --lean --no-cpu --workload $wl 2>/dev/null | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print('$wl', 'fuse_append=$f', 'value', j['value'], 'e2e', j['e2e']['value'], 'e2e_ms', j['e2e']['ms_per_step'], j['config']['num_splits'])"
done; done 2>&1 | tee -a gpurun_out/e2e_ab.txt
