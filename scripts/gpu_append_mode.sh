# usage (under gpurun): bash scripts/gpu_append_mode.sh -- whole GPU suite (append tests run in both store modes), then the bench step per mode
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) > gpurun_out/mode_pytest.log; cat gpurun_out/mode_pytest.log
for i in 1 2; do for m in 1 2; do
B2_APPEND_MODE=$m timeout 100 python bench.py --steps 300 --warmup 20 --lean --no-cpu 2>/dev/null | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('append_mode=$m', 'value', j['value'], 'ms', j['ms_per_step'], 'kern', r['kernel_ms'], 'min', r['kernel_ms_min'])"
done; done 2>&1 | tee gpurun_out/append_mode_ab.txt
