# usage (under gpurun --gpus 2): bash scripts/gpu_n2_diag.sh -- is a slow N=2 line the box (host cores, topology, tenants) or the code?
mkdir -p gpurun_out
{
  echo "nproc $(nproc)  cpu.max $(cat /sys/fs/cgroup/cpu.max 2>/dev/null)  loadavg $(cat /proc/loadavg)"
  python - <<'PY'
import os
print("sched_getaffinity", len(os.sched_getaffinity(0)), "cpu_count", os.cpu_count())
PY
  nvidia-smi topo -m 2>&1 | head -12
  nvidia-smi --query-gpu=index,utilization.gpu,memory.used,clocks.sm,pstate --format=csv
  nvidia-smi --query-compute-apps=pid,used_memory --format=csv
} > gpurun_out/n2_diag_env.txt 2>&1
cat gpurun_out/n2_diag_env.txt
run() {  # tag, env...
  tag=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus 2 --steps 20 --warmup 5 --lean --no-cpu 2> gpurun_out/n2_diag_$tag.err | tail -1 > gpurun_out/n2_diag_$tag.json
  python - <<PY
import json
try:
    j=json.load(open('gpurun_out/n2_diag_$tag.json'))
    r=j['roofline']; e=j['e2e']
    print('$tag', 'value', j['value'], 'ms', j['ms_per_step'], 'kern', r['kernel_ms'], 'kern_min', r['kernel_ms_min'], 'e2e', e['value'], 'e2e_ms', e['ms_per_step'])
except Exception as ex:
    print('$tag', 'ERR', ex)
PY
}
run pdl_on  B2_DECODE_PDL=1
run pdl_off B2_DECODE_PDL=0
run pdl_on2 B2_DECODE_PDL=1
echo "loadavg after: $(cat /proc/loadavg)"
# one rank alone on each GPU, to see whether one of the two devices is the slow one
for g in 0 1; do
  CUDA_VISIBLE_DEVICES=$g timeout 200 python bench.py --steps 20 --warmup 5 --lean --no-cpu 2>/dev/null | tail -1 > gpurun_out/n2_diag_gpu$g.json
  python - <<PY
import json
try:
    j=json.load(open('gpurun_out/n2_diag_gpu$g.json'))
    r=j['roofline']; e=j['e2e']
    print('gpu$g alone', 'value', j['value'], 'kern', r['kernel_ms'], 'kern_min', r['kernel_ms_min'], 'e2e', e['value'], 'e2e_ms', e['ms_per_step'])
except Exception as ex:
    print('gpu$g', 'ERR', ex)
PY
done
