# r03v: block-level occlusion test — parity (incl. on/off equality at full size), then A/B
cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r03v_parity.txt 2>&1; tail -4 gpurun_out/r03v_parity.txt
for bc in 0 1; do
export UGU_B200_BLOCKCULL=$bc
timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r03v_bench_bc$bc.json 2> gpurun_out/r03v_bench_bc$bc.err
timeout 600 python tools/bench_configs.py c5 c4 --no-cpu > gpurun_out/r03v_configs_bc$bc.jsonl 2> gpurun_out/r03v_configs_bc$bc.err
python - $bc <<'PY'
import json,sys
bc=sys.argv[1]
d=json.loads(open(f'gpurun_out/r03v_bench_bc{bc}.json').read().strip().splitlines()[-1])
print('blockcull',bc,'c2', round(d['ms_per_step']*1e3,1), d['ms_per_step_spread']['median'], d['stage_us'])
for l in open(f'gpurun_out/r03v_configs_bc{bc}.jsonl'):
    if l.startswith('{'):
        c=json.loads(l); print('blockcull',bc,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
done
timeout 900 python -m pytest tests/test_raytrace.py -m gpu -x -q -k "ordered or full_size" > gpurun_out/r03v_rt.txt 2>&1; tail -2 gpurun_out/r03v_rt.txt
