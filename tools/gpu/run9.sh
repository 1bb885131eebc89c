# r02j: two-pass occlusion ordering of raster_kernel — parity, then A/B against the single pass
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02j_parity.txt 2>&1; tail -5 gpurun_out/r02j_parity.txt
for tp in 0 1; do
UGU_B200_TWOPASS=$tp timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r02j_bench_tp$tp.json 2> gpurun_out/r02j_bench_tp$tp.err
UGU_B200_TWOPASS=$tp timeout 600 python tools/bench_configs.py c5 c4 c3 --no-cpu > gpurun_out/r02j_configs_tp$tp.jsonl 2> gpurun_out/r02j_configs_tp$tp.err
done
python - <<'PY'
import json
for tp in (0,1):
    d=json.loads(open(f'gpurun_out/r02j_bench_tp{tp}.json').read().strip().splitlines()[-1])
    print('tp',tp,'c2', round(d['ms_per_step']*1e3,1), d['stage_us'])
    for l in open(f'gpurun_out/r02j_configs_tp{tp}.jsonl'):
        if l.startswith('{'):
            c=json.loads(l); print('tp',tp,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
