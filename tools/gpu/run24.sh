# r03c: raster_views_kernel (a CTA keeps its block for all views of a launch set) — parity of the batch paths, A/B
cd $GRAFT_REPO_ROOT
UGU_B200_VIEWS_LOOP=1 timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_bands.py tests/test_multiview.py -m gpu -x -q -k "batch or multiview or config5 or config3 or guard" > gpurun_out/r03c_parity.txt 2>&1; tail -3 gpurun_out/r03c_parity.txt
for vl in 0 1; do
export UGU_B200_VIEWS_LOOP=$vl
timeout 600 python tools/bench_configs.py c5 c3 c1 --no-cpu > gpurun_out/r03c_configs_vl$vl.jsonl 2> gpurun_out/r03c_configs_vl$vl.err
python - $vl <<'PY'
import json,sys
vl=sys.argv[1]
for l in open(f'gpurun_out/r03c_configs_vl{vl}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('views_loop',vl,c['config'][:8], round(c['ms_per_view']*1e3,1), 'batch', round(c['batch_ms_per_view']['median']*1e3,2), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
