# r02n: single-launch occlusion ordering (block lists from the vertex kernel) — parity, then A/B
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02n_parity.txt 2>&1; tail -5 gpurun_out/r02n_parity.txt
for cfg in "0 5" "1 5" "1 6"; do set -- $cfg
export UGU_B200_TWOPASS=$1 UGU_B200_RASTER_MINB=$2
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r02n_bench_tp$1_b$2.json 2> gpurun_out/r02n_bench_tp$1_b$2.err
timeout 600 python tools/bench_configs.py c5 c4 --no-cpu > gpurun_out/r02n_configs_tp$1_b$2.jsonl 2> gpurun_out/r02n_configs_tp$1_b$2.err
python - $1 $2 <<'PY'
import json,sys
tp,b=sys.argv[1:3]
d=json.loads(open(f'gpurun_out/r02n_bench_tp{tp}_b{b}.json').read().strip().splitlines()[-1])
print('tp',tp,'minb',b,'c2', round(d['ms_per_step']*1e3,1), d['stage_us'])
for l in open(f'gpurun_out/r02n_configs_tp{tp}_b{b}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('tp',tp,'minb',b,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
export UGU_B200_TWOPASS=1 UGU_B200_RASTER_MINB=5
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-multiview"
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum --clock-control none --kernel-name-base demangled -k regex:raster_kernel -c 4 --csv \
  --log-file gpurun_out/r02n_launches.csv $CMD > gpurun_out/r02n_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02n_launches.csv')) if len(r)>10]
for r in rows[-3:]: print(r[4][:40], r[-3], r[-1])
PY
