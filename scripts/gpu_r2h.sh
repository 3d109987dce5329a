# ragged batches after the r2 decode changes + final ncu capture of the rope ring kernel + element-wise op numbers
tag=${1:-r2h}
sed -i 's/--no-cpu --ragged/--no-cpu --lean --ragged/' scripts/gpu_ragged.sh
bash scripts/gpu_ragged.sh > gpurun_out/ragged_$tag.txt 2>&1; cat gpurun_out/ragged_$tag.txt | tail -9
python scripts/ops_once.py > /dev/null 2>&1 && WRITE_K=0 ncu --set full --clock-control none --import-source on -k regex:'rope_and_cache' -s 2 -c 1 -f -o gpurun_out/prof_${tag}_rope_ring python scripts/ops_once.py > gpurun_out/ncu_${tag}_rope.log 2>&1
timeout 300 python scripts/bench_model.py --only ops > gpurun_out/ops_$tag.jsonl 2>&1; cut -c1-150 gpurun_out/ops_$tag.jsonl
