cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02e_gpu_tests.txt 2>&1; tail -5 gpurun_out/r02e_gpu_tests.txt
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r02e_bench.json 2> gpurun_out/r02e_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r02e_bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['stage_us'], d['e2e']['ms_per_step'])"
timeout 600 python tools/bench_configs.py c1 c3 c5 c4 --no-cpu > gpurun_out/r02e_configs.jsonl 2> gpurun_out/r02e_configs.err
python - <<'PY'
import json
for l in open('gpurun_out/r02e_configs.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print(d['config'][:8], round(d['ms_per_view']*1e3,1), d['stage_us'], 'batch', round(d['batch_ms_per_view']['median']*1e3,1), d['batch_ms_per_view']['stage_us_per_view'])
PY
