# r03b: tile jobs folded into the resolve kernel (no binned_kernel launch) — parity, then A/B
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_bands.py -m gpu -x -q > gpurun_out/r03b_parity.txt 2>&1; tail -3 gpurun_out/r03b_parity.txt
for fb in 0 1; do
export UGU_B200_FOLD_BINNED=$fb
timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r03b_bench_fold$fb.json 2> gpurun_out/r03b_bench_fold$fb.err
timeout 600 python tools/bench_configs.py c5 --no-cpu > gpurun_out/r03b_configs_fold$fb.jsonl 2> gpurun_out/r03b_configs_fold$fb.err
python - $fb <<'PY'
import json,sys
fb=sys.argv[1]
d=json.loads(open(f'gpurun_out/r03b_bench_fold{fb}.json').read().strip().splitlines()[-1])
print('fold',fb,'c2', round(d['ms_per_step']*1e3,1), d['ms_per_step_spread']['median'], d['stage_us'])
for l in open(f'gpurun_out/r03b_configs_fold{fb}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('fold',fb,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
