cd $GRAFT_REPO_ROOT
for mb in 0 1; do
export UGU_B200_RESOLVE_MINB3=$mb
timeout 600 python tools/bench_configs.py c1 c3 c5 --no-cpu > gpurun_out/r03h_configs_mb$mb.jsonl 2> gpurun_out/r03h_configs_mb$mb.err
python - $mb <<'PY'
import json,sys
mb=sys.argv[1]
for l in open(f'gpurun_out/r03h_configs_mb{mb}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('resolve3',mb,c['config'][:8], round(c['ms_per_view']*1e3,1), 'batch', round(c['batch_ms_per_view']['median']*1e3,2), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
