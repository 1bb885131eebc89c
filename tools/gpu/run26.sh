# r03f: bulk L2 prefetches (raster streams from the vertex kernel, visibility buffer from the resolve kernel) — A/B
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "golden or full_size or batch" > gpurun_out/r03f_parity.txt 2>&1; tail -3 gpurun_out/r03f_parity.txt
for pf in 0 1; do
export UGU_B200_PREFETCH=$pf
timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r03f_bench_pf$pf.json 2> gpurun_out/r03f_bench_pf$pf.err
timeout 600 python tools/bench_configs.py c5 c4 --no-cpu > gpurun_out/r03f_configs_pf$pf.jsonl 2> gpurun_out/r03f_configs_pf$pf.err
python - $pf <<'PY'
import json,sys
fb=sys.argv[1]
d=json.loads(open(f'gpurun_out/r03f_bench_pf{fb}.json').read().strip().splitlines()[-1])
print('prefetch',fb,'c2', round(d['ms_per_step']*1e3,1), d['ms_per_step_spread']['median'], d['stage_us'])
for l in open(f'gpurun_out/r03f_configs_pf{fb}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('prefetch',fb,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
