# usage (under gpurun --gpus 2): bash scripts/gpu_tp_leftover.sh <tag> -- do the tensor-parallel tests leave a process on a GPU?
tag=${1:-a}
mkdir -p gpurun_out
snap() {
  echo "== $1 ($(date +%s)) loadavg $(cat /proc/loadavg)"
  nvidia-smi --query-compute-apps=pid,used_memory --format=csv,noheader
  nvidia-smi --query-gpu=index,utilization.gpu,memory.used --format=csv,noheader
  ps -eo pid,ppid,etimes,pcpu,stat,args | grep -E "b200attn|tp_check|torch.distributed|pytest" | grep -v grep | cut -c1-200
}
{
  snap before
  for t in test_tp2_engine_matches_oracle test_server_tensor_parallel_size_2; do
    t0=$(date +%s)
    timeout 400 python -m pytest "tests/test_gpu_tp.py::$t" -m gpu -x -q 2>&1 | tail -3
    echo "$t took $(( $(date +%s) - t0 )) s"
    snap "right after $t"
    sleep 8
    snap "8 s after $t"
  done
} > gpurun_out/tp_leftover_$tag.txt 2>&1
cat gpurun_out/tp_leftover_$tag.txt
