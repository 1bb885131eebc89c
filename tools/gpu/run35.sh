cd $GRAFT_REPO_ROOT
for pc in 1 2 4; do
export UGU_B200_PER_CTA=$pc
timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r03w_bench_pc$pc.json 2> gpurun_out/r03w_bench_pc$pc.err
timeout 600 python tools/bench_configs.py c5 --no-cpu > gpurun_out/r03w_configs_pc$pc.jsonl 2> gpurun_out/r03w_configs_pc$pc.err
python - $pc <<'PY'
import json,sys
pc=sys.argv[1]
d=json.loads(open(f'gpurun_out/r03w_bench_pc{pc}.json').read().strip().splitlines()[-1])
print('per_cta',pc,'c2', round(d['ms_per_step']*1e3,1), d['stage_us'])
for l in open(f'gpurun_out/r03w_configs_pc{pc}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('per_cta',pc,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
