# r02p: occlusion ordering variants: 0 = off, 1 = ordered single launch, 2 = two launches (CTA-level classification)
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02p_parity.txt 2>&1; tail -3 gpurun_out/r02p_parity.txt
UGU_B200_TWOPASS=2 timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "full_size or debris or soup" > gpurun_out/r02p_parity_tp2.txt 2>&1; tail -3 gpurun_out/r02p_parity_tp2.txt
for tp in 0 1 2; do
export UGU_B200_TWOPASS=$tp
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r02p_bench_tp$tp.json 2> gpurun_out/r02p_bench_tp$tp.err
timeout 600 python tools/bench_configs.py c5 c4 --no-cpu > gpurun_out/r02p_configs_tp$tp.jsonl 2> gpurun_out/r02p_configs_tp$tp.err
python - $tp <<'PY'
import json,sys
tp=sys.argv[1]
d=json.loads(open(f'gpurun_out/r02p_bench_tp{tp}.json').read().strip().splitlines()[-1])
print('tp',tp,'c2', round(d['ms_per_step']*1e3,1), d['stage_us'])
for l in open(f'gpurun_out/r02p_configs_tp{tp}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('tp',tp,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
export UGU_B200_TWOPASS=2
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-multiview"
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum --clock-control none --kernel-name-base demangled -k regex:raster_kernel -c 4 --csv \
  --log-file gpurun_out/r02p_launches.csv $CMD > gpurun_out/r02p_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02p_launches.csv')) if len(r)>10]
for r in rows[-6:]: print(r[4][:40], r[-3], r[-1])
PY
