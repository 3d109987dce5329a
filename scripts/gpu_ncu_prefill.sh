# usage: bash scripts/gpu_ncu_prefill.sh <tag>  (under gpurun, 1 GPU): --set full capture of the prefill kernel
tag=${1:-x}
python scripts/prefill_once.py 3 > gpurun_out/plain_prefill_$tag.log 2>&1 || { tail -5 gpurun_out/plain_prefill_$tag.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:prefill -s 2 -c 1 -f -o gpurun_out/prof_prefill_$tag python scripts/prefill_once.py 3 > gpurun_out/ncu_prefill_$tag.log 2>&1
tail -2 gpurun_out/ncu_prefill_$tag.log
