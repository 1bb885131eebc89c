# r03d: covered-sample compaction across the tile jobs of a triangle — parity + configs 1/3 (+ncu counts)
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_bands.py tests/test_gbuffer.py -m gpu -x -q > gpurun_out/r03d_parity.txt 2>&1; tail -3 gpurun_out/r03d_parity.txt
timeout 600 python tools/bench_configs.py c1 c3 --no-cpu > gpurun_out/r03d_configs.jsonl 2> gpurun_out/r03d_configs.err
python - <<'PY'
import json
for l in open('gpurun_out/r03d_configs.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print(c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,2), c['batch_ms_per_view']['stage_us_per_view'])
PY
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none --kernel-name-base demangled -k regex:"binned_kernel" -s 4 -c 2 --csv --log-file gpurun_out/r03d_c3_binned.csv python tools/bench_configs.py c3 --no-cpu > /dev/null 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r03d_c3_binned.csv')) if len(r)>10]
for r in rows[1:]: print(r[4][:34], r[-3][:40], r[-1])
PY
