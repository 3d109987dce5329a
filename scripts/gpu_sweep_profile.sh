set -x
for s in 1 2 4 8 16 32; do timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu --splits $s 2>&1 | tail -1 >> gpurun_out/sweep_r1.jsonl; done
timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu --workload llama3-8b_b128_ctx8k_fp8kv 2>&1 | tail -1 >> gpurun_out/sweep_r1.jsonl
timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu --workload llama3-70b_b256_ctx8k_bf16kv 2>&1 | tail -1 >> gpurun_out/sweep_r1.jsonl
timeout 300 python bench.py --steps 300 --warmup 20 --no-cpu --workload oracle_b4_ctx512 2>&1 | tail -1 >> gpurun_out/sweep_r1.jsonl
python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/plain_r1.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_decode|reshape_and_cache|merge_splits' -c 400 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/ncu_launch_r1.log 2>&1
python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/plain_r1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:paged_decode_kernel -s 8 -c 2 -f -o gpurun_out/prof_r1 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/ncu_full_r1.log 2>&1
ls -la gpurun_out
