for B in 37 74 64 111; do for s in 0 1; do
timeout 300 python bench.py --steps 300 --warmup 20 --no-cpu --batch $B --splits $s 2>&1 | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('B',j['config']['global_batch'],'splits',j['config']['num_splits'],'kern_ms',r['kernel_ms'],'min',r['kernel_ms_min'],'ach',r['achieved'])"
done; done
python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/plain_r1.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_decode|merge_' -s 20 -c 20 --csv --log-file gpurun_out/launches_r1d.csv python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/ncu_launch_r1d.log 2>&1
python - <<PY
import csv
for r in csv.DictReader(l for l in open('gpurun_out/launches_r1d.csv') if l.startswith('"')):
    print(r['Kernel Name'][:60], r['Grid Size'], r['Metric Value'])
PY
