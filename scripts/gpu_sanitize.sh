# usage: bash scripts/gpu_sanitize.sh <tag>  (under gpurun, 1 GPU): compute-sanitizer over scripts/sanitize_cases.py
tag=${1:-san}
python scripts/sanitize_cases.py > gpurun_out/sanitize_plain_$tag.log 2>&1 || { tail -20 gpurun_out/sanitize_plain_$tag.log; exit 1; }
tail -3 gpurun_out/sanitize_plain_$tag.log
for tool in memcheck racecheck initcheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 40 python scripts/sanitize_cases.py > gpurun_out/sanitize_${tool}_$tag.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|cases ok|cases done|Error|hazard" gpurun_out/sanitize_${tool}_$tag.log | sort | uniq -c | head -12
done
# launch list of the bench command, decode-path kernels only (the r2c list was filled by the problem set-up kernels)
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --lean"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_decode|reshape_and_cache|merge_|zero_counter' -c 400 --csv \
    --log-file gpurun_out/launches_$tag.csv $CMD > gpurun_out/ncu_launch_$tag.log 2>&1
# NVML NVLink counters: do they exist on this box?
python - <<'PY' > gpurun_out/nvlink_probe.txt 2>&1
import pynvml
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
for fid in (138, 139):
    for scope in (0xFFFFFFFF, 0):
        try:
            v = pynvml.nvmlDeviceGetFieldValues(h, [(fid, scope)])
            print(fid, hex(scope), v[0].nvmlReturn, v[0].valueType, v[0].value.ullVal)
        except Exception as e:
            print(fid, hex(scope), 'EXC', e)
PY
cat gpurun_out/nvlink_probe.txt; nvidia-smi nvlink -gt d -i 0 2>&1 | head -8
