# usage: bash scripts/gpu_e2e.sh <tag> <tp>   (under gpurun --gpus tp): BASELINE configs[4] end to end through the
# evaluation-Job template, served at --tensor-parallel-size tp
tag=${1:-e2e}; tp=${2:-1}
timeout 1500 python scripts/run_e2e.py --tensor-parallel-size $tp --tag ${tag}_tp$tp > gpurun_out/e2e_${tag}_tp$tp.log 2>&1
echo "e2e rc=$?"; grep -v "^b200attn_\|^#" gpurun_out/e2e_${tag}_tp$tp.log | tail -25 | cut -c1-600
