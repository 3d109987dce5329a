# usage: bash scripts/gpu_tp.sh <tag> <ngpus> [legs]   (under gpurun --gpus N): comm kernel tests, then the TP check
tag=${1:-tp}; n=${2:-2}; legs=${3:-parity,exchange,decode}
(timeout 600 python -m pytest tests/test_gpu_comm.py -q -x 2>&1 | tail -25) > gpurun_out/pytest_comm_$tag.log 2>&1
tail -6 gpurun_out/pytest_comm_$tag.log
nvidia-smi topo -m > gpurun_out/topo_$tag.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    scripts/tp_check.py --legs $legs > gpurun_out/tp_check_${tag}.json 2> gpurun_out/tp_check_${tag}.err
echo "tp_check rc=$?"; tail -c 1500 gpurun_out/tp_check_${tag}.err; tail -c 3000 gpurun_out/tp_check_${tag}.json
