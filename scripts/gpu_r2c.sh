# usage: bash scripts/gpu_r2c.sh <tag>  (under gpurun, 1 GPU): GPU tests, driver-style bench line, reference arm,
# ncu launch list of the bench command, one --set full capture per decode kernel variant (+ config 4)
tag=${1:-r2c}
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -40) > gpurun_out/pytest_$tag.log 2>&1
tail -5 gpurun_out/pytest_$tag.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_${tag}_k20.json 2> gpurun_out/bench_${tag}_k20.err
echo "bench k20 rc=$?"; tail -c 600 gpurun_out/bench_${tag}_k20.err
timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_${tag}_ref.json 2>&1
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --lean"
$CMD > gpurun_out/plain_$tag.log 2>&1 || { tail -5 gpurun_out/plain_$tag.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches_$tag.csv $CMD > gpurun_out/ncu_launch_$tag.log 2>&1
for v in "llama3-8b_b64_ctx4k_bf16kv 1" "llama3-8b_b64_ctx4k_bf16kv 0" "llama3-8b_b128_ctx8k_fp8kv 0" "llama3-70b_b256_ctx8k_bf16kv 1"; do
  set -- $v
  C2="python bench.py --steps 8 --warmup 3 --no-cpu --lean --workload $1 --splits $2"
  $C2 > gpurun_out/plain_${tag}_$1_s$2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'paged_decode_(split|dynamic)_kernel' -s 6 -c 1 -f \
      -o gpurun_out/prof_${tag}_$1_s$2 $C2 > gpurun_out/ncu_full_${tag}_$1_s$2.log 2>&1
done
ls -la gpurun_out
