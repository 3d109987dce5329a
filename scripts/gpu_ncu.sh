# usage: bash scripts/gpu_ncu.sh <tag>   (run under gpurun, 1 GPU)
# 1) plain run must exit 0, 2) launch list of the same bench command, 3) one --set full capture per kernel variant
tag=${1:-x}
set -x
CMD="python bench.py --steps 20 --warmup 5 --no-cpu"
$CMD > gpurun_out/plain_$tag.log 2>&1 || { tail -5 gpurun_out/plain_$tag.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_decode|reshape_and_cache|merge_|zero_counter' -c 400 --csv \
    --log-file gpurun_out/launches_$tag.csv $CMD > gpurun_out/ncu_launch_$tag.log 2>&1
for v in "llama3-8b_b64_ctx4k_bf16kv 1" "llama3-8b_b64_ctx4k_bf16kv 0" "llama3-8b_b128_ctx8k_fp8kv 1" "llama3-8b_b128_ctx8k_fp8kv 0"; do
  set -- $v
  C2="python bench.py --steps 8 --warmup 3 --no-cpu --workload $1 --splits $2"
  $C2 > gpurun_out/plain_${tag}_$1_s$2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'paged_decode_(split|dynamic)_kernel' -s 6 -c 1 -f \
      -o gpurun_out/prof_${tag}_$1_s$2 $C2 > gpurun_out/ncu_full_${tag}_$1_s$2.log 2>&1
done
ls -la gpurun_out
