# usage (under gpurun): bash scripts/gpu_r2_final.sh -- the round's closing check on the final code:
# the whole GPU suite as the driver runs it, smoke(), the driver-style bench line, the reference arm, prefill numbers.
mkdir -p gpurun_out
t0=$(date +%s)
(timeout 1100 python -m pytest tests -m gpu -x -q --durations=12 2>&1 | tail -45) > gpurun_out/final_pytest.log 2>&1
echo "pytest wall $(( $(date +%s) - t0 )) s" >> gpurun_out/final_pytest.log
tail -5 gpurun_out/final_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; tail -1 gpurun_out/final_smoke.log
t0=$(date +%s)
timeout 500 python bench.py --steps 20 --warmup 5 > gpurun_out/final_bench_n1.log 2>&1
echo "bench wall $(( $(date +%s) - t0 )) s"
tail -1 gpurun_out/final_bench_n1.log > gpurun_out/final_bench_n1.json
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 > gpurun_out/final_bench_reference.json
timeout 300 python scripts/bench_model.py --only prefill 2>&1 | tail -12 > gpurun_out/final_prefill.jsonl
head -c 1500 gpurun_out/final_bench_n1.json; echo
cat gpurun_out/final_prefill.jsonl | cut -c1-300
