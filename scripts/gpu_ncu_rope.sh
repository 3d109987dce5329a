tag=${1:-rope}
python scripts/ops_once.py > gpurun_out/ops_once_$tag.log 2>&1 || { tail gpurun_out/ops_once_$tag.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'rope_and_cache' -s 2 -c 1 -f -o gpurun_out/prof_${tag}_stream python scripts/ops_once.py > gpurun_out/ncu_${tag}_stream.log 2>&1
B2_ROPE_IMPL=token ncu --set full --clock-control none --import-source on -k regex:'rope_and_cache' -s 2 -c 1 -f -o gpurun_out/prof_${tag}_token python scripts/ops_once.py > gpurun_out/ncu_${tag}_token.log 2>&1
SLOTS=random ncu --set full --clock-control none --import-source on -k regex:'rope_and_cache' -s 2 -c 1 -f -o gpurun_out/prof_${tag}_stream_randslots python scripts/ops_once.py > gpurun_out/ncu_${tag}_stream_rand.log 2>&1
timeout 300 python scripts/bench_model.py --only ops 2>&1 | grep rope | cut -c1-300
ls gpurun_out/*.ncu-rep | tail -4
