n=${1:-8}
for v in "0" "1"; do
B2_COMM_DEBUG=1 B2_COMM_WEAK_ST=$v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    scripts/tp_check.py --legs exchange > gpurun_out/tp_dbg_${n}_weak$v.json 2> gpurun_out/tp_dbg_${n}_weak$v.err
echo "rc=$?"; tail -c 2500 gpurun_out/tp_dbg_${n}_weak$v.json
done
