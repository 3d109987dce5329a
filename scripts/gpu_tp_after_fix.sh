# usage (under gpurun --gpus 2): tensor-parallel + server tests, a process snapshot, then the driver's N=2 bench line right behind them
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_tp.py tests/test_gpu_server.py tests/test_gpu_comm.py -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tp_after_fix_pytest.log
cat gpurun_out/tp_after_fix_pytest.log
{
  echo "== right after the tests: loadavg $(cat /proc/loadavg)"
  nvidia-smi --query-compute-apps=pid,used_memory --format=csv,noheader
  nvidia-smi --query-gpu=index,utilization.gpu,memory.used --format=csv,noheader
  ps -eo pid,ppid,etimes,pcpu,stat,args | grep -E "b200attn|tp_check" | grep -v grep | cut -c1-160
  echo "== end of snapshot"
} > gpurun_out/tp_after_fix_snapshot.txt 2>&1
cat gpurun_out/tp_after_fix_snapshot.txt
bash scripts/gpu_bench_n.sh after_fix 2 20
