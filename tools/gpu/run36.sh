cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_gbuffer.py -m gpu -x -q > gpurun_out/r03y_parity.txt 2>&1; tail -2 gpurun_out/r03y_parity.txt
timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r03y_bench.json 2> gpurun_out/r03y_bench.err
timeout 600 python tools/bench_configs.py c5 c4 c3 --no-cpu > gpurun_out/r03y_configs.jsonl 2> gpurun_out/r03y_configs.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r03y_bench.json').read().strip().splitlines()[-1])
print('c2', round(d['ms_per_step']*1e3,1), d['ms_per_step_spread']['median'], d['stage_us'])
for l in open('gpurun_out/r03y_configs.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print(c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
