# r02r: per-tile triangle lists (tile_kernel) vs the tile-job queue (binned_kernel) on the large-triangle configs
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_bands.py tests/test_gbuffer.py -m gpu -x -q > gpurun_out/r02r_parity.txt 2>&1; tail -5 gpurun_out/r02r_parity.txt
for tl in 0 1; do
export UGU_B200_TILES=$tl
timeout 600 python tools/bench_configs.py c1 c3 --no-cpu > gpurun_out/r02r_configs_tiles$tl.jsonl 2> gpurun_out/r02r_configs_tiles$tl.err
python - $tl <<'PY'
import json,sys
tl=sys.argv[1]
for l in open(f'gpurun_out/r02r_configs_tiles{tl}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('tiles',tl,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
